/*
 * ubench_pipes.cu -- per-SM issue rates of the instructions the stand kernels lean on
 * (F2F.F64.F32, F2F.F32.F64, DADD, DFMA, FFMA, LDS.128), measured on the box so that the
 * single-read stand design is sized against numbers, not guesses.
 * build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_pipes tools/ubench_pipes.cu
 * prints thread-instructions per clock per SM for each op (all SMs busy, 1024 threads/SM).
 */
#include <cstdio>
#include <cuda_runtime.h>

template <int OP> __global__ void __launch_bounds__ (1024)
k (float *out, float seed, int iters, long long *cycles)
{
  float f[8];
  double d[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    f[i] = seed + i + threadIdx.x;
    d[i] = (double) seed * (i + 1) + threadIdx.x;
  }
  const long long t0 = clock64 ();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (OP == 0) { d[i] = (double) f[i]; f[i] = __int_as_float (__float_as_int (f[i]) ^ (int) __double2loint (d[i])); }   /* F2F.F64.F32 + LOP */
      if (OP == 1) { f[i] = (float) d[i]; d[i] = __hiloint2double (__double2hiint (d[i]), __float_as_int (f[i])); }        /* F2F.F32.F64 */
      if (OP == 2) d[i] = __dadd_rn (d[i], 1.25);
      if (OP == 3) d[i] = fma (d[i], 1.0000001, 0.5);
      if (OP == 4) f[i] = fmaf (f[i], 1.0000001f, 0.5f);
      if (OP == 5) d[i] = __dmul_rn (d[i], 1.0000001);
    }
  }
  const long long t1 = clock64 ();
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 8; i++)
    acc += f[i] + (float) d[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0)
    cycles[blockIdx.x] = t1 - t0;
}

template <int OP> static void
run (const char *name, int sms)
{
  float *out;
  long long *cyc;
  cudaMalloc (&out, sizeof (float) * 1024 * sms);
  cudaMalloc (&cyc, sizeof (long long) * sms);
  const int iters = 4096;
  k<OP><<<sms, 1024>>> (out, 1.5f, 16, cyc);
  k<OP><<<sms, 1024>>> (out, 1.5f, iters, cyc);
  cudaDeviceSynchronize ();
  long long h[256];
  cudaMemcpy (h, cyc, sizeof (long long) * sms, cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < sms; i++)
    avg += (double) h[i];
  avg /= sms;
  printf ("%-14s %8.2f thread-ops/clk/SM  (%.0f cycles)\n", name, 1024.0 * 8 * iters / avg, avg);
  cudaFree (out);
  cudaFree (cyc);
}

int
main ()
{
  cudaDeviceProp p;
  cudaGetDeviceProperties (&p, 0);
  printf ("%s, %d SMs\n", p.name, p.multiProcessorCount);
  run<0> ("F2F.F64.F32", p.multiProcessorCount);
  run<1> ("F2F.F32.F64", p.multiProcessorCount);
  run<2> ("DADD", p.multiProcessorCount);
  run<3> ("DFMA", p.multiProcessorCount);
  run<5> ("DMUL", p.multiProcessorCount);
  run<4> ("FFMA", p.multiProcessorCount);
  return 0;
}
