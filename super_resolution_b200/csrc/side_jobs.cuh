// side_jobs.cuh -- the two per-sample "finish" steps of a HAB (ECA gate of the CAB branch, dynamic depthwise kernel of the
// ESC branch) as device functions, so that they can ride as EXTRA CTAs of a large kernel that is launched anyway
// (tc_conv3.cu: cab_conv1 carries dyn_finish; tc_plk.cu: the 13x13 kernel carries eca_finish) instead of costing a launch
// of their own on an otherwise idle GPU (5.8 us x 36 per forward in the ncu launch list of round 2).
#pragma once
#include "common.cuh"

namespace hb {

// Sum of `n` per-tile partials `stride` floats apart, in index order (deterministic), with 16 loads in flight:
// these finish kernels are one small block per sample and nothing but the latency of this dependent chain.
__device__ __forceinline__ float chunk_sum(const float* __restrict__ p, int n, int stride) {
  float s = 0.f;
  for (int k0 = 0; k0 < n; k0 += 16) {
    float t[16];
#pragma unroll
    for (int j = 0; j < 16; j++) t[j] = k0 + j < n ? __ldg(p + (size_t)(k0 + j) * stride) : 0.f;
#pragma unroll
    for (int j = 0; j < 16; j++) s += t[j];
  }
  return s;
}

// ECA gate: sigmoid(conv1d_k5(mean over HW)) across the channel axis, zero padded
// ends, no bias (hat_arch.py:66-77).  One block per sample.
__device__ __forceinline__ void eca_finish_body(int b, const float* __restrict__ partial, int nchunk, int C, float inv_hw,
                                                const float* __restrict__ w5, float* __restrict__ scale, float* mean) {
  for (int c = threadIdx.x; c < C; c += blockDim.x) mean[c] = chunk_sum(partial + (size_t)b * nchunk * C + c, nchunk, C) * inv_hw;
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f;
#pragma unroll
    for (int j = 0; j < 5; j++) {
      int cc = c + j - 2;
      if (cc >= 0 && cc < C) a = fmaf(w5[j], mean[cc], a);
    }
    scale[(size_t)b * C + c] = sigmoid_f(a);
  }
}

// ESC dynamic depthwise kernel per sample (esc_arch.py:95-100, :121):
// dyn = W2 . gelu(W1 . GAP(x[:, :pdim]) + b1) + b2, reshaped (pdim, 3, 3).
__device__ __forceinline__ void dyn_finish_body(int b, const float* __restrict__ partial, int nchunk, int pdim, float inv_hw,
                                                const float* __restrict__ w1, const float* __restrict__ b1,
                                                const float* __restrict__ w2, const float* __restrict__ b2,
                                                float* __restrict__ dyn, float* mean, float* hid) {
  const int half = pdim / 2;
  for (int c = threadIdx.x; c < pdim; c += blockDim.x)
    mean[c] = chunk_sum(partial + (size_t)b * nchunk * pdim + c, nchunk, pdim) * inv_hw;
  __syncthreads();
  for (int o = threadIdx.x; o < half; o += blockDim.x) {
    float a = b1[o];
    for (int i = 0; i < pdim; i++) a = fmaf(w1[o * pdim + i], mean[i], a);
    hid[o] = gelu_f(a);
  }
  __syncthreads();
  for (int j = threadIdx.x; j < pdim * 9; j += blockDim.x) {
    float a = b2[j];
    for (int i = 0; i < half; i++) a = fmaf(w2[j * half + i], hid[i], a);
    dyn[(size_t)b * pdim * 9 + j] = a;
  }
}


// Runs `job` for sample `b` with the whole CTA; `scratch` >= (max(C, 64) + 32) floats of shared memory.
__device__ __forceinline__ void run_side_job(const SideJob& job, int b, float* scratch) {
  if (b >= job.B) return;
  if (job.kind == 1) eca_finish_body(b, job.partial, job.nchunk, job.C, job.inv_hw, job.w1, job.out, scratch);
  else if (job.kind == 2)
    dyn_finish_body(b, job.partial, job.nchunk, job.C, job.inv_hw, job.w1, job.b1, job.w2, job.b2, job.out, scratch, scratch + 64);
}

}  // namespace hb
