// Debug micro-benchmark (not part of the library): issue cost of the jet phase-1 inner loop
// and of its ingredients on one SM sub-partition.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mix_bench tools/mix_bench.cu
// Prints cycles per loop iteration (one "slot x tile" = 16 pixels per lane) for 1..6 warps
// per scheduler.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

typedef unsigned long long f2;
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ f2 f2_pack(float lo, float hi) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void f2_unpack(f2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f2 f2_fma(f2 a, f2 b, f2 c) { f2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f2 f2_add(f2 a, f2 b) { f2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f2 f2_mul(f2 a, f2 b) { f2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// mode 0: packed body as in jet2.cu   1: same without MUFU   2: scalar FFMA body with MUFU
// mode 3: MUFU only                   4: FFMA2 only (32 per iteration)   5: FFMA only (64)
// mode 6: second-order recurrence along rows (2 MUFU per column pair... see below)
template <int MODE>
__global__ void __launch_bounds__(1024, 1) mix_kernel(const float4* coef, int iters, float* out, long long* cycles) {
  __shared__ float4 sc[64];
  if (threadIdx.x < 64) sc[threadIdx.x] = coef[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const float c0f = (float)(2 * lane), r0f = 3.f;
  f2 acc2[8];
  float acc[16];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc2[i] = 0ull;
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  const long long t0 = clock64();
  float4 q = sc[0], ct = sc[1];
  for (int it = 0; it < iters; ++it) {
    float4 qn, ctn;
    if (MODE == 7) {
      const int kn = (it + 1) & 15;
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(qn.x), "=f"(qn.y), "=f"(qn.z), "=f"(qn.w) : "r"((unsigned)__cvta_generic_to_shared(&sc[2 * kn])));
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(ctn.x), "=f"(ctn.y), "=f"(ctn.z), "=f"(ctn.w) : "r"((unsigned)__cvta_generic_to_shared(&sc[2 * kn + 1])));
    } else {
      const int k = it & 15;
      q = sc[2 * k]; ct = sc[2 * k + 1];
    }
    const float d0 = (c0f - ct.z) - ct.w;
    const float dr0 = (r0f - ct.x) - ct.y;
    if (MODE == 0 || MODE == 1 || MODE == 7) {
      const f2 d = f2_pack(d0, d0 + 1.f);
      const f2 Q = f2_mul(f2_pack(q.y, q.y), d);
      const f2 Pq = f2_mul(f2_mul(f2_pack(q.x, q.x), d), d);
      const f2 amp2 = f2_pack(q.w, q.w);
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        const f2 drp = f2_pack(dr0 + (float)i, dr0 + (float)(i + 1));
        const f2 Cqp = f2_mul(f2_mul(f2_pack(q.z, q.z), drp), drp);
        float dra, drb, Cqa, Cqb, ta, tb;
        f2_unpack(drp, dra, drb);
        f2_unpack(Cqp, Cqa, Cqb);
        f2_unpack(f2_fma(f2_pack(dra, dra), Q, f2_add(Pq, f2_pack(Cqa, Cqa))), ta, tb);
        if (MODE != 1) { ta = ex2(ta); tb = ex2(tb); }
        acc2[i] = f2_fma(amp2, f2_pack(ta, tb), acc2[i]);
        f2_unpack(f2_fma(f2_pack(drb, drb), Q, f2_add(Pq, f2_pack(Cqb, Cqb))), ta, tb);
        if (MODE != 1) { ta = ex2(ta); tb = ex2(tb); }
        acc2[i + 1] = f2_fma(amp2, f2_pack(ta, tb), acc2[i + 1]);
      }
    } else if (MODE == 2) {
      const float d1 = d0 + 1.f;
      const float Q0 = q.y * d0, Q1 = q.y * d1, P0 = q.x * d0 * d0, P1 = q.x * d1 * d1;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float dr = dr0 + (float)i;
        const float Cq = q.z * dr * dr;
        acc[2 * i] = fmaf(q.w, ex2(fmaf(dr, Q0, P0 + Cq)), acc[2 * i]);
        acc[2 * i + 1] = fmaf(q.w, ex2(fmaf(dr, Q1, P1 + Cq)), acc[2 * i + 1]);
      }
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = ex2(acc[i] + d0);
    } else if (MODE == 4) {
      const f2 a = f2_pack(d0, dr0), b = f2_pack(q.x, q.y);
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc2[i] = f2_fma(a, acc2[i], b);
    } else if (MODE == 5) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fmaf(d0, acc[i], q.x);
    } else if (MODE == 6) {
      // rows in two blocks of four: E0 and W per column and block on the SFU (4 MUFU per
      // block and column pair instead of 8), powers of W and the per-slot constants
      // K[i] = 2^(C i^2) on the FMA pipe
      const f2 d = f2_pack(d0, d0 + 1.f);
      const f2 Q = f2_mul(f2_pack(q.y, q.y), d);
      const f2 Pq = f2_mul(f2_mul(f2_pack(q.x, q.x), d), d);
      const f2 amp2 = f2_pack(q.w, q.w);
      const f2 K1 = f2_pack(ct.x, ct.x), K2 = f2_pack(ct.y, ct.y), K3 = f2_pack(ct.z, ct.z);
#pragma unroll
      for (int b = 0; b < 8; b += 4) {
        const float dr = dr0 + (float)b;
        const float Cq = q.z * dr * dr;
        float ta, tb, wa, wb;
        f2_unpack(f2_fma(f2_pack(dr, dr), Q, f2_add(Pq, f2_pack(Cq, Cq))), ta, tb);
        const float twoC = 2.f * q.z * dr;
        f2_unpack(f2_add(Q, f2_pack(twoC, twoC)), wa, wb);
        const f2 E0 = f2_mul(amp2, f2_pack(ex2(ta), ex2(tb)));
        const f2 W1 = f2_pack(ex2(wa), ex2(wb));
        const f2 W2 = f2_mul(W1, W1), W3 = f2_mul(W2, W1);
        acc2[b] = f2_add(acc2[b], E0);
        acc2[b + 1] = f2_fma(E0, f2_mul(W1, K1), acc2[b + 1]);
        acc2[b + 2] = f2_fma(E0, f2_mul(W2, K2), acc2[b + 2]);
        acc2[b + 3] = f2_fma(E0, f2_mul(W3, K3), acc2[b + 3]);
      }
    }
    if (MODE == 7) { q = qn; ct = ctn; }
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { float a, b; f2_unpack(acc2[i], a, b); s += a + b; }
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, const float4* d_coef, float* d_out, long long* d_cyc) {
  const int iters = 4096;
  printf("%-34s", name);
  for (int wps = 1; wps <= 8; wps *= 2) {  // warps per scheduler
    const int threads = 32 * 4 * wps;
    mix_kernel<MODE><<<148, threads>>>(d_coef, iters, d_out, d_cyc);
    mix_kernel<MODE><<<148, threads>>>(d_coef, iters, d_out, d_cyc);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double m = 0;
    for (int i = 0; i < 148; ++i) m += (double)h[i];
    m /= 148.0;
    printf("  %dw: %7.1f (%6.1f/warp-iter)", wps, m / iters, m / iters / wps);
  }
  printf("\n");
}

int main() {
  float4 h_coef[64];
  for (int i = 0; i < 64; ++i) h_coef[i] = make_float4(-0.01f - 0.001f * i, 0.002f, -0.012f, 0.7f);
  for (int i = 1; i < 64; i += 2) h_coef[i] = make_float4(0.99f, 0.98f, 0.97f, 3.f);
  float4* d_coef; float* d_out; long long* d_cyc;
  cudaMalloc(&d_coef, sizeof(h_coef));
  cudaMalloc(&d_out, 148 * 1024 * 4);
  cudaMalloc(&d_cyc, 148 * 8);
  cudaMemcpy(d_coef, h_coef, sizeof(h_coef), cudaMemcpyHostToDevice);
  printf("cycles per iteration (16 pixels per lane) with N warps per scheduler\n");
  run<0>("0 packed body + 16 MUFU", d_coef, d_out, d_cyc);
  run<1>("1 packed body, no MUFU", d_coef, d_out, d_cyc);
  run<2>("2 scalar body + 16 MUFU", d_coef, d_out, d_cyc);
  run<3>("3 16 MUFU + 16 FADD", d_coef, d_out, d_cyc);
  run<4>("4 32 FFMA2", d_coef, d_out, d_cyc);
  run<5>("5 64 FFMA", d_coef, d_out, d_cyc);
  run<6>("6 4-row blocks, 8 MUFU", d_coef, d_out, d_cyc);
  run<7>("7 packed body, coefs prefetched", d_coef, d_out, d_cyc);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
