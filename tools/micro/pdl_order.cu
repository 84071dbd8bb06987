// What does a stream operation that follows a kernel which has executed griddepcontrol.launch_dependents wait for?
// A: triggers at its top, spins ~200 us, then writes flag = 1.  Then, in the same stream:
//   case 0: plain kernel B reads flag                      (expected 1)
//   case 1: cudaMemsetAsync(other) ; plain kernel B        (expected 1)
//   case 2: kernel B launched with the PDL attribute, B calls griddepcontrol.wait first   (expected 1)
//   case 3: cudaMemsetAsync(flag2 = 0xff..) ; B launched with the PDL attribute, waits, reads flag2  (does B wait for the memset?)
//   case 4: A ; PDL kernel C (trigger at top, wait, no-op) ; cudaMemsetAsync(other) ; plain B reads flag
#include <cstdio>
#include <cuda_runtime.h>
__global__ void A(volatile int* flag, long long spin)
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const long long t0 = clock64();
    while (clock64() - t0 < spin) { }
    if (threadIdx.x == 0 && blockIdx.x == 0) *flag = 1;
}
__global__ void C()
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
// launched with the PDL attribute: waits, (optionally triggers), spins, then writes flag = 1
__global__ void D(volatile int* flag, long long spin, int trigger)
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (trigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const long long t0 = clock64();
    while (clock64() - t0 < spin) { }
    if (threadIdx.x == 0 && blockIdx.x == 0) *flag = 1;
}
__global__ void B(const volatile int* flag, int* out, int wait)
{
    if (wait) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0 && blockIdx.x == 0) *out = *flag;
}
template <typename... K, typename... T>
void launch(bool pdl, void (*k)(K...), int grid, cudaStream_t s, T... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = 64; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, k, static_cast<K>(args)...);
}
int main()
{
    cudaStream_t s; cudaStreamCreate(&s);
    int *flag, *flag2, *out, *other; cudaMalloc(&flag, 4); cudaMalloc(&flag2, 4); cudaMalloc(&out, 4); cudaMalloc(&other, 1 << 20);
    for (int c = 0; c < 9; c++) {
        int bad = 0;
        for (int rep = 0; rep < 200; rep++) {
            cudaMemsetAsync(flag, 0, 4, s); cudaMemsetAsync(flag2, 0, 4, s); cudaMemsetAsync(out, 0xff, 4, s);
            cudaStreamSynchronize(s);
            launch(false, A, 148, s, (volatile int*)flag, 400000LL);
            if (c == 0) launch(false, B, 1, s, (const volatile int*)flag, out, 0);
            if (c == 1) { cudaMemsetAsync(other, 0, 1 << 20, s); launch(false, B, 1, s, (const volatile int*)flag, out, 0); }
            if (c == 2) launch(true, B, 1, s, (const volatile int*)flag, out, 1);
            if (c == 3) { cudaMemsetAsync(flag2, 0x01, 4, s); launch(true, B, 1, s, (const volatile int*)flag2, out, 1); }
            if (c == 4) { launch(true, C, 148, s); cudaMemsetAsync(other, 0, 1 << 20, s); launch(false, B, 1, s, (const volatile int*)flag, out, 0); }
            // 5/6: A ; D (PDL-launched, long, writes flag=1 at its end; 6: D also triggers) ; memset(flag = 0x02..): the memset must win
            if (c == 5 || c == 6) { launch(true, D, 148, s, (volatile int*)flag, 400000LL, c == 6); cudaMemsetAsync(flag, 0x02, 4, s); launch(false, B, 1, s, (const volatile int*)flag, out, 0); }
            // 7/8: A ; D (PDL-launched, long) ; plain B reads flag (must see 1)
            if (c == 7 || c == 8) { launch(true, D, 148, s, (volatile int*)flag, 400000LL, c == 8); launch(false, B, 1, s, (const volatile int*)flag, out, 0); }
            int h = -1;
            cudaMemcpyAsync(&h, out, 4, cudaMemcpyDeviceToHost, s);
            cudaStreamSynchronize(s);
            const int want = (c == 3) ? 0x01010101 : ((c == 5 || c == 6) ? 0x02020202 : 1);
            if (h != want) bad++;
        }
        printf("case %d: %d of 200 runs saw the wrong value (%s)\n", c, bad, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
