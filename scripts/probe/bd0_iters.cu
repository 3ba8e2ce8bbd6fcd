#include <cstdio>
#include <cfloat>
#include <cmath>
__device__ int bd0_iters(double np, double* out) {
  if (fabs(1.0 - np) < 0.1 * (1.0 + np)) {
    double v = (1.0 - np) / (1.0 + np);
    double s = (1.0 - np) * v;
    if (fabs(s) < DBL_MIN) { *out = s; return 0; }
    double ej = 2 * v;
    v = v * v;
    for (int j = 1; j < 1000; j++) {
      ej *= v;
      double s1 = s + ej / (double)((j << 1) + 1);
      if (s1 == s) { *out = s1; return j; }
      s = s1;
    }
    return 1000;
  }
  *out = log(1.0 / np) + np - 1.0;
  return -1;
}
__global__ void k(int n, int* hist, double* worst) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double eta = -12.0 + 24.0 * i / n;
  double p = 1. / (1. + exp(-eta));
  double q = 1 - p, r;
  int it = -2;
  if (p < 0.1 && p != 0) it = bd0_iters(q, &r);
  else if (q < 0.1 && q != 0) it = bd0_iters(p, &r);
  if (it >= 0) atomicAdd(&hist[it > 40 ? 40 : it], 1);
  if (it >= 30) worst[0] = eta;
}
int main() {
  int n = 1 << 24, *hist; double* worst;
  cudaMallocManaged(&hist, 41 * sizeof(int)); cudaMallocManaged(&worst, 8);
  for (int i = 0; i < 41; i++) hist[i] = 0; worst[0] = 0;
  k<<<(n + 255) / 256, 256>>>(n, hist, worst);
  cudaDeviceSynchronize();
  for (int i = 0; i < 41; i++) if (hist[i]) printf("iters %d: %d\n", i, hist[i]);
  printf("worst eta %.6f err %s\n", worst[0], cudaGetErrorString(cudaGetLastError()));
}
