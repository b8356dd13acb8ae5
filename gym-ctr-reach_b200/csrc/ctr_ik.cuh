// ctr_ik.cuh -- finite-difference Jacobian and damped-least-squares inverse kinematics, included by
// ctr_kernels.cu.
//
// Algorithm of the reference's baseline controller, src/jacobian_controller.py:19-73 (dls_ik_position_only) and
// envs/CTR_Python/CTR_Model.py:251-262 (jac): forward differences with eps = 1e-4 on the 6 joints (7 kinematics solves per
// iteration), dq = J^T (J J^T + lam I)^-1 e with lam = 0.25, q += dq, stop once the error BEFORE that update was below 1e-3 m,
// at most 500 iterations.  The reference script differentiates the BVP model (CTR_Model, out of scope) and imports
// modules that are not in its tree, so it cannot be run: this is that algorithm over the environment's forward kinematics
// (Model.forward_kinematics = ctr_fk), and its parity is pinned to the restatement in oracle/ik_oracle.py only.
// The kinematics solves are k_fk launches on a 7n-column joint block; the kernels here are the small HBM-bound glue.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ctr {

// q f64 [6][n] -> block f64 [6][7n]: column 7 i + k holds q_i with joint k - 1 advanced by eps (k = 0: unperturbed).  The 7 solves
// of one configuration are neighbours, so they share a warp of k_fk (10 solves in lock-step): they have the same section table and
// take the same steps, which is what a lock-step warp wants (column k n + i, the first layout, spread them over 7 warps with 9
// unrelated solves each: 392 us against 335 us per 16384 Jacobians).  system_out i32 [7n] (or null) gets the same layout.
// Frozen (converged) configurations are still written: their columns cost a solve but keep the launch shape fixed.
__global__ void k_ik_perturb(const double* __restrict__ q, const int32_t* __restrict__ system, int64_t n, double eps,
                             double* __restrict__ block, int32_t* __restrict__ system_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) v[c] = q[(int64_t)c * n + i];
    const int sy = system ? system[i] : 0;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
#pragma unroll
        for (int c = 0; c < 6; ++c)
            block[(int64_t)c * 7 * n + 7 * i + k] = (k == c + 1) ? __dadd_rn(v[c], eps) : v[c];   // CTR_Model.py:257
        if (system_out) system_out[7 * i + k] = sy;
    }
}

// tips f64 [3][7n] -> J f64 [18][n], row-major 3 x 6: J[r][c] = (tip_c+1[r] - tip_0[r]) / eps   (CTR_Model.py:259-260)
// CM (configuration major): solve k of configuration i is column 7 i + k (k_ik_perturb); otherwise k n + i (the variant blocks of
// the BVP back-end's launches, csrc/ctr_bvp.cuh)
template <bool CM>
__device__ __forceinline__ void fd_jacobian(const double* __restrict__ tips, int64_t n, int64_t i, double eps, double (&P)[3],
                                            double (&J)[3][6]) {
#pragma unroll
    for (int r = 0; r < 3; ++r) P[r] = tips[(int64_t)r * 7 * n + (CM ? 7 * i : i)];
#pragma unroll
    for (int c = 0; c < 6; ++c)
#pragma unroll
        for (int r = 0; r < 3; ++r)
            J[r][c] = __ddiv_rn(__dsub_rn(tips[(int64_t)r * 7 * n + (CM ? 7 * i + (c + 1) : (int64_t)(c + 1) * n + i)], P[r]), eps);
}

__global__ void k_ik_jacobian(const double* __restrict__ tips, int64_t n, double eps, double* __restrict__ Jout,
                              double* __restrict__ tip0) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double P[3], J[3][6];
    fd_jacobian<true>(tips, n, i, eps, P, J);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        if (tip0) tip0[(int64_t)r * n + i] = P[r];
#pragma unroll
        for (int c = 0; c < 6; ++c) Jout[(int64_t)(r * 6 + c) * n + i] = J[r][c];
    }
}

// One DLS iteration for every configuration that has not converged (jacobian_controller.py:41-53):
//   e = target - P;  q += J^T (J J^T + lam I)^-1 e;  converged once ||e|| < tol (the update of that iteration is applied).
// err[i] = ||e|| of the last iteration run, iters[i] = iterations run, active[0] counts the configurations still iterating.
template <bool CM>
__global__ void k_ik_dls_update(const double* __restrict__ tips, const double* __restrict__ target, int64_t n, double eps,
                                double lam, double tol, double* __restrict__ q, double* __restrict__ err,
                                int32_t* __restrict__ iters, uint8_t* __restrict__ frozen, int32_t* __restrict__ active) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || frozen[i]) return;
    double P[3], J[3][6], e[3];
    fd_jacobian<CM>(tips, n, i, eps, P, J);
#pragma unroll
    for (int r = 0; r < 3; ++r) e[r] = __dsub_rn(target[(int64_t)r * n + i], P[r]);
    double A[3][3];                                                   // J J^T + lam I (symmetric)
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            double s = 0.0;
#pragma unroll
            for (int c = 0; c < 6; ++c) s = fma(J[a][c], J[b][c], s);
            A[a][b] = (a == b) ? s + lam : s;
        }
    // x = A^-1 e by the adjugate (A is symmetric positive definite for lam > 0)
    const double c00 = A[1][1] * A[2][2] - A[1][2] * A[2][1], c01 = A[1][2] * A[2][0] - A[1][0] * A[2][2],
                 c02 = A[1][0] * A[2][1] - A[1][1] * A[2][0];
    const double det = A[0][0] * c00 + A[0][1] * c01 + A[0][2] * c02;
    const double c11 = A[0][0] * A[2][2] - A[0][2] * A[2][0], c12 = A[0][1] * A[2][0] - A[0][0] * A[2][1],
                 c22 = A[0][0] * A[1][1] - A[0][1] * A[1][0];
    const double x0 = (c00 * e[0] + c01 * e[1] + c02 * e[2]) / det, x1 = (c01 * e[0] + c11 * e[1] + c12 * e[2]) / det,
                 x2 = (c02 * e[0] + c12 * e[1] + c22 * e[2]) / det;
#pragma unroll
    for (int c = 0; c < 6; ++c) q[(int64_t)c * n + i] += J[0][c] * x0 + J[1][c] * x1 + J[2][c] * x2;   // q += J^T x
    const double en = sqrt((e[0] * e[0] + e[1] * e[1]) + e[2] * e[2]);
    err[i] = en;
    iters[i] += 1;
    if (en < tol) frozen[i] = 1;                                      // jacobian_controller.py:50-51: break after the update
    else atomicAdd(active, 1);
}

}  // namespace ctr
