// Pose optimiser step on the device: Adam on the 6-vector twist followed by the retraction
//   T <- Exp(step) * T
// as ONE single-thread kernel.  CF-3DGS does this with a torch.optim.Adam parameter group and lietorch's SE3 retraction
// (scene/gaussian_model_cf.py:303-373: ~45 tiny PyTorch kernels per iteration when written with tensor ops); in a
// pose-fit loop of 300 iterations per frame pair those launches are a tenth of the iteration once the rasterizer is
// fast.  The twist gradient is what gsr_backward / gsr_backward_rigid write to dL_dpose (translation first).
//
//   t += 1;  m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2
//   step = -lr * (m / (1 - b1^t)) / (sqrt(v / (1 - b2^t)) + eps)            (torch.optim.Adam, no weight decay)
//   Exp((rho, w)) = [ R  V rho ; 0 1 ],  R = I + A K + B K^2,  V = I + B K + C K^2,  K = hat(w),
//   A = sin(th)/th, B = (1 - cos th)/th^2, C = (th - sin th)/th^3   (series below 1e-4 rad)
// Everything is evaluated in double by one thread (a few hundred flops), so the float32 pose carries no error from
// the update itself.
#include "../../include/gsr.h"
#include "gsr_kernels.h"
#include "gsr_launch.cuh"

namespace gsr {
void set_last_error(const char* msg);

namespace {

__global__ void pose_adam_step_kernel(const float* __restrict__ grad, float* __restrict__ state, float* __restrict__ T,
                                      float lr, float b1, float b2, float eps)
{
    pdl_wait();
    pdl_trigger();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    // state: m[6], v[6], t
    const double t = (double)state[12] + 1.0;
    state[12] = (float)t;
    const double c1 = 1.0 - pow((double)b1, t), c2 = 1.0 - pow((double)b2, t);
    double xi[6];
    for (int k = 0; k < 6; k++) {
        const double g = (double)grad[k];
        const double m = (double)b1 * (double)state[k] + (1.0 - (double)b1) * g;
        const double v = (double)b2 * (double)state[6 + k] + (1.0 - (double)b2) * g * g;
        state[k] = (float)m;
        state[6 + k] = (float)v;
        xi[k] = -(double)lr * (m / c1) / (sqrt(v / c2) + (double)eps);
    }
    const double wx = xi[3], wy = xi[4], wz = xi[5];
    const double th2 = wx * wx + wy * wy + wz * wz, th = sqrt(th2);
    double A, B, C;
    if (th2 < 1e-8) {
        A = 1.0 - th2 / 6.0; B = 0.5 - th2 / 24.0; C = 1.0 / 6.0 - th2 / 120.0;
    } else {
        A = sin(th) / th; B = (1.0 - cos(th)) / th2; C = (th - sin(th)) / (th2 * th);
    }
    const double K[3][3] = {{0.0, -wz, wy}, {wz, 0.0, -wx}, {-wy, wx, 0.0}};
    double K2[3][3], R[3][3], V[3][3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) K2[i][j] = K[i][0] * K[0][j] + K[i][1] * K[1][j] + K[i][2] * K[2][j];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            const double I = i == j ? 1.0 : 0.0;
            R[i][j] = I + A * K[i][j] + B * K2[i][j];
            V[i][j] = I + B * K[i][j] + C * K2[i][j];
        }
    double E[4][4] = {{0}};
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) E[i][j] = R[i][j];
        E[i][3] = V[i][0] * xi[0] + V[i][1] * xi[1] + V[i][2] * xi[2];
    }
    E[3][3] = 1.0;
    double Tn[4][4];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            Tn[i][j] = E[i][0] * (double)T[j] + E[i][1] * (double)T[4 + j] + E[i][2] * (double)T[8 + j] + E[i][3] * (double)T[12 + j];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) T[4 * i + j] = (float)Tn[i][j];
}

}  // namespace
}  // namespace gsr

using namespace gsr;

extern "C" int gsr_pose_adam_step(const float* grad, float* state, float* T, float lr, float beta1, float beta2, float eps,
                                  void* stream)
{
    if (!grad || !state || !T) { set_last_error("gsr_pose_adam_step: grad, state and T are required"); return GSR_ERR_INVALID; }
    launch_pdl(pose_adam_step_kernel, 1, 32, 0, (cudaStream_t)stream, grad, state, T, lr, beta1, beta2, eps);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); return GSR_ERR_CUDA; }
    return GSR_OK;
}
