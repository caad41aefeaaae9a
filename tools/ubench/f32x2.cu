// Micro-benchmark: issue rate of scalar vs packed (f32x2) FP32 instructions on sm_100a, alone and mixed
// with shared-memory loads / integer ops.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o f32x2 f32x2.cu
#include <cuda_runtime.h>
#include <cstdio>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c){ u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b){ u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c){ float r; asm("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ float add1(float a, float b){ float r; asm("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }

constexpr int ITERS = 4096;
constexpr int NACC = 16;

// mode 0: FFMA x16/iter; 1: FFMA2 x16; 2: FADD x16; 3: FADD2 x16; 4: FFMA2 x16 + LDS.64 x8; 5: FFMA x16 + LDS.64 x8;
// 6: FFMA2 x16 with swapped operand; 7: FFMA2 x8 + FFMA x16 (interleaved); 8: FFMA2 x16 + IADD x16; 9: FFMA x16 + IADD x16
template <int MODE>
__global__ void __launch_bounds__(256) bench(const float* in, float* out, long long* cycles) {
  __shared__ float2 sm[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = make_float2(in[i & 63], in[(i + 7) & 63]);
  __syncthreads();
  float a[NACC], b[NACC];
  u64 p[NACC];
  int ia[NACC];
  for (int i = 0; i < NACC; ++i) { a[i] = in[i]; b[i] = in[i + 16]; p[i] = pk(a[i], b[i]); ia[i] = threadIdx.x + i; }
  const float m = in[40], c = in[41];
  const u64 m2 = pk(m, in[42]), c2 = pk(c, in[43]);
  float2 acc = make_float2(0.f, 0.f);
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      if (MODE == 0 || MODE == 5 || MODE == 9) a[i] = fma1(a[i], m, c);
      if (MODE == 1 || MODE == 4 || MODE == 8) p[i] = fma2(p[i], m2, c2);
      if (MODE == 2) a[i] = add1(a[i], c);
      if (MODE == 3) p[i] = add2(p[i], c2);
      if (MODE == 6) { float x, y; upk(p[i], x, y); p[i] = fma2(pk(y, -x), pk(m, m), c2); }
      if (MODE == 7) { a[i] = fma1(a[i], m, c); if (i & 1) p[i] = fma2(p[i], m2, c2); }
      if ((MODE == 4 || MODE == 5) && (i & 1)) { float2 v = sm[(threadIdx.x + i * 32 + it) & 1023]; acc.x += 0.f * v.x; acc.y = v.y; }
      if (MODE == 8 || MODE == 9) asm volatile("add.s32 %0, %0, %1;" : "+r"(ia[i]) : "r"(it));
    }
  }
  long long t1 = clock64();
  float s = acc.x + acc.y;
  u64 tot = p[0];
  for (int i = 1; i < NACC; ++i) tot = add2(tot, p[i]);
  { float x, y; upk(tot, x, y); s += x + y; }
  for (int i = 0; i < NACC; ++i) s += a[i] + b[i] + ia[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int per_iter_fp, int warps_per_sm, const float* in, float* out, long long* cyc) {
  int threads = 256, blocks = 148 * warps_per_sm / 8;
  bench<MODE><<<blocks, threads>>>(in, out, cyc);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  bench<MODE><<<blocks, threads>>>(in, out, cyc);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[2048]; cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < blocks; ++i) avg += h[i]; avg /= blocks;
  // warp-instructions of the named kind per SMSP per clock
  double warps_per_smsp = warps_per_sm / 4.0;
  double ipc = (double)ITERS * per_iter_fp * warps_per_smsp / avg;
  printf("%-44s warps/SM %2d  cycles %9.0f  fp-instr/clk/SMSP %.3f  (%.3f ms)  err=%s\n", name, warps_per_sm, avg, ipc, ms,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  float h[64]; for (int i = 0; i < 64; ++i) h[i] = 1.0f + i * 1e-3f;
  float *in, *out; long long* cyc;
  cudaMalloc(&in, sizeof(h)); cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
  cudaMalloc(&out, 4 * 256 * 148 * 8); cudaMalloc(&cyc, 8 * 2048);
  for (int w : {8, 16, 32}) {
    run<0>("FFMA x16", 16, w, in, out, cyc);
    run<1>("FFMA2 x16", 16, w, in, out, cyc);
    run<2>("FADD x16", 16, w, in, out, cyc);
    run<3>("FADD2 x16", 16, w, in, out, cyc);
    run<6>("FFMA2 x16 swapped-operand", 16, w, in, out, cyc);
    run<7>("FFMA x16 + FFMA2 x8 (count 24)", 24, w, in, out, cyc);
    run<5>("FFMA x16 + LDS.64 x8 (count fp only)", 16, w, in, out, cyc);
    run<4>("FFMA2 x16 + LDS.64 x8 (count fp only)", 16, w, in, out, cyc);
    run<9>("FFMA x16 + IADD x16 (count fp only)", 16, w, in, out, cyc);
    run<8>("FFMA2 x16 + IADD x16 (count fp only)", 16, w, in, out, cyc);
  }
  return 0;
}
