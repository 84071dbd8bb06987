// Programmatic dependent launch (sm_90+): every kernel of the forward / backward chain is launched with the
// "programmatic stream serialization" attribute and begins with pdl_wait().  The hardware may then schedule a kernel's
// blocks while the kernel before it in the stream is still draining -- block scheduling, shared-memory set-up and
// address arithmetic overlap the predecessor's tail -- and pdl_wait() holds each thread until every earlier kernel has
// finished and its memory is visible.  The chain is ~35 kernels of 3-50 us each, most of them a single wave, so the
// ramp / drain at every boundary is a measurable share of the step.
//
// Rules the kernels follow: nothing is read from or written to global memory before pdl_wait() (an earlier kernel may
// still be using it); pdl_wait() is executed by every thread that goes on to touch global memory; pdl_trigger() only
// lets the NEXT kernel start its own prologue early and carries no memory guarantee.
//
// pdl_wait() comes BEFORE pdl_trigger() in every kernel.  With the trigger first a kernel lets its successor start
// while it is itself still waiting, three or more kernels of the chain become resident at once, and on B200 (driver
// 580) that order produced wrong results (a status kernel read the instance counter before the kernel two places up
// the chain had finished adding to it; tools/micro/pdl_order.cu holds the orderings that were checked and do hold).
//
// Rule for the host side: the attribute is only used when the operation before the launch on that stream is one of
// these kernels.  Every memset, copy or event record breaks the chain (chain_break()), and so does every entry into the
// library: the first kernel after any of them is launched the ordinary, fully serialised way.
#pragma once
#include <cuda_runtime.h>

namespace gsr {

#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

bool& pdl_chain_flag();                                  // per host thread: was the previous stream operation one of our kernels?
inline void chain_break() { pdl_chain_flag() = false; }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args)
{
    bool& chained = pdl_chain_flag();
    const bool overlap = chained;
    chained = true;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = overlap ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace gsr
