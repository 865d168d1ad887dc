#include <cstdio>
#include <cuda_runtime.h>
#define UG_NDOF 12
__global__ void k(const float* Min, float* out, float* piv) {
  __shared__ float sM[12][12];
  int lane = threadIdx.x;
  for (int i = lane; i < 144; i += 32) (&sM[0][0])[i] = Min[i];
  __syncwarp();
  bool isd = lane < 12;
  {
    float L[UG_NDOF][UG_NDOF];
#pragma unroll
    for (int i = 0; i < UG_NDOF; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) L[i][j] = sM[i][j];
#pragma unroll
    for (int k = 0; k < UG_NDOF; k++) {
      float dk = L[k][k];
#pragma unroll
      for (int t = 0; t < k; t++) dk -= L[k][t] * L[k][t];
      float inv = rsqrtf(dk);
      if (lane == 0) piv[k] = dk;
      L[k][k] = inv;
#pragma unroll
      for (int i = k + 1; i < UG_NDOF; i++) {
        float v = L[i][k];
#pragma unroll
        for (int t = 0; t < k; t++) v -= L[i][t] * L[k][t];
        L[i][k] = v * inv;
      }
    }
    float x[UG_NDOF];
#pragma unroll
    for (int i = 0; i < UG_NDOF; i++) {
      float v = (i == lane) ? 1.0f : 0.0f;
#pragma unroll
      for (int t = 0; t < i; t++) v -= L[i][t] * x[t];
      x[i] = v * L[i][i];
    }
#pragma unroll
    for (int i = UG_NDOF - 1; i >= 0; i--) {
      float v = x[i];
#pragma unroll
      for (int t = i + 1; t < UG_NDOF; t++) v -= L[t][i] * x[t];
      x[i] = v * L[i][i];
    }
    __syncwarp();
    if (isd) {
#pragma unroll
      for (int i = 0; i < UG_NDOF; i++) out[i * 12 + lane] = x[i];
    }
  }
}
int main() {
  float M[144];
  for (int i = 0; i < 12; i++) for (int j = 0; j < 12; j++) M[i*12+j] = (i == j) ? 2.0f + i : 0.1f / (1 + i + j);
  float *dM, *dO, *dP; cudaMalloc(&dM, 576); cudaMalloc(&dO, 576); cudaMalloc(&dP, 48);
  cudaMemcpy(dM, M, 576, cudaMemcpyHostToDevice);
  k<<<1, 32>>>(dM, dO, dP);
  float O[144], P[12]; cudaMemcpy(O, dO, 576, cudaMemcpyDeviceToHost); cudaMemcpy(P, dP, 48, cudaMemcpyDeviceToHost);
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
  for (int i = 0; i < 12; i++) printf("piv %g minv_ii %g\n", P[i], O[i*12+i]);
  // check M*Minv
  double e = 0; for (int i = 0; i < 12; i++) for (int j = 0; j < 12; j++) { double s = 0; for (int t = 0; t < 12; t++) s += (double)M[i*12+t] * O[t*12+j]; e = fmax(e, fabs(s - (i == j))); }
  printf("max |M Minv - I| = %g\n", e);
}
