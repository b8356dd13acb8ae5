// accuracy of the MUFU.RCP64H seed and of one / two Newton steps over the range of the RMS-norm scales
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>
__global__ void k(double* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    double worst0 = 0, worst1 = 0, worst2 = 0;
    for (int k = i; k < n; k += gridDim.x * blockDim.x) {
        // log-uniform in [1e-7, 1e1] with a scrambled mantissa
        unsigned h = k * 2654435761u;
        double x = exp2(-24.0 + 28.0 * (double)(h >> 8) / 16777216.0) * (1.0 + (double)(h & 0xff) / 977.0);
        double r;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
        double e0 = fabs(fma(-x, r, 1.0));
        double e = fma(-x, r, 1.0);
        double r1 = fma(r, e, r);
        double e1 = fabs(fma(-x, r1, 1.0));
        e = fma(-x, r1, 1.0);
        double r2 = fma(r1, e, r1);
        double e2 = fabs(fma(-x, r2, 1.0));
        worst0 = fmax(worst0, e0); worst1 = fmax(worst1, e1); worst2 = fmax(worst2, e2);
    }
    out[3 * i] = worst0; out[3 * i + 1] = worst1; out[3 * i + 2] = worst2;
}
int main() {
    const int T = 256 * 592; double* d; cudaMalloc(&d, T * 3 * sizeof(double));
    k<<<592, 256>>>(d, 1 << 28);
    double* h = new double[T * 3]; cudaMemcpy(h, d, T * 3 * sizeof(double), cudaMemcpyDeviceToHost);
    double w[3] = {0, 0, 0};
    for (int i = 0; i < T; ++i) for (int j = 0; j < 3; ++j) w[j] = fmax(w[j], h[3 * i + j]);
    printf("max |1 - x*r|: seed %.3e (2^%.1f)  1 Newton %.3e (2^%.1f)  2 Newton %.3e\n", w[0], log2(w[0]), w[1], log2(w[1]), w[2]);
    return 0;
}
