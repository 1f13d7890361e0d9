// attn_simt.cu -- OCAB overlapping-window attention core on CUDA cores (fp32 math).
// Reference: OCAB.forward, hat_arch.py:354-388.  One block per (window, head), one
// thread per query.  Keys/values are fetched straight from the token tensor by 2-D
// coordinates (no unfold tensor); positions outside the image are ZERO vectors that
// still take softmax mass (nn.Unfold pads after projection, hat_arch.py:296-297).
// The relative-position bias comes from the compact (heads, (ws+ows-1)^2) table built
// by pack_ocab_bias from the wrapped index buffer (hat_arch.py:378-382).
#include "kernels.cuh"

namespace hb {

constexpr int KCHUNK_ROWS = 8;   // key rows staged per pass

template <typename T, int DMAX>
__global__ void __launch_bounds__(256) attn_simt_kernel(const T* __restrict__ qkv, int pitch, int H, int W, int C,
                                                        int heads, int ws, int ows, const float* __restrict__ bias,
                                                        T* __restrict__ out, int out_pitch, int out_zero_to, int pad,
                                                        const float* __restrict__ kadd) {
  extern __shared__ float smem[];
  const int d = C / heads;
  const int dp = (d + 3) & ~3;
  const int Tn = ws + ows - 1;
  float* sbias = smem;                          // Tn*Tn
  float* sk = sbias + ((Tn * Tn + 3) & ~3);     // KCHUNK_ROWS*ows*dp
  float* sv = sk + KCHUNK_ROWS * ows * dp;

  const int nwx = W / ws, nwy = H / ws;
  const int win = blockIdx.x, head = blockIdx.y;
  const int b = win / (nwx * nwy);
  const int wrem = win % (nwx * nwy);
  const int wy = wrem / nwx, wx = wrem % nwx;
  const float* ka = kadd ? kadd + (size_t)win * ows * ows : nullptr;    // HATX focus bias / top-k pruning
  const int tid = threadIdx.x;
  const int nq = ws * ws;
  const bool active = tid < nq;
  const int qy = tid / ws, qx = tid % ws;
  const size_t qtok = ((size_t)b * H + wy * ws + qy) * W + wx * ws + qx;

  for (int i = tid; i < Tn * Tn; i += blockDim.x) sbias[i] = bias[(size_t)head * Tn * Tn + i];

  float q[DMAX], acc[DMAX];
#pragma unroll
  for (int i = 0; i < DMAX; i++) {
    q[i] = (active && i < d) ? to_f(qkv[qtok * pitch + head * d + i]) : 0.f;
    acc[i] = 0.f;
  }
  float mx = -INFINITY, l = 0.f;

  for (int ky0 = 0; ky0 < ows; ky0 += KCHUNK_ROWS) {
    const int rows = min(KCHUNK_ROWS, ows - ky0);
    __syncthreads();
    for (int e = tid; e < rows * ows * dp; e += blockDim.x) {
      int i = e % dp;
      int kk = e / dp;
      int ky = ky0 + kk / ows, kx = kk % ows;
      int yy = wy * ws - pad + ky, xx = wx * ws - pad + kx;
      float kv = 0.f, vv = 0.f;
      if (i < d && yy >= 0 && yy < H && xx >= 0 && xx < W) {
        size_t tok = ((size_t)b * H + yy) * W + xx;
        kv = to_f(qkv[tok * pitch + C + head * d + i]);
        vv = to_f(qkv[tok * pitch + 2 * C + head * d + i]);
      }
      sk[e] = kv;
      sv[e] = vv;
    }
    __syncthreads();
    if (!active) continue;
    for (int kr = 0; kr < rows; kr++) {
      const int ky = ky0 + kr;
      const float* brow = sbias + (ky - qy + ws - 1) * Tn + (ws - 1 - qx);
      for (int kx = 0; kx < ows; kx++) {
        const float extra = ka ? ka[ky * ows + kx] : 0.f;
        if (extra == -INFINITY) continue;       // pruned key: exp(-1e4 + bias - max) is exactly 0 (hatx_arch.py:437-441)
        const float* kp = sk + (kr * ows + kx) * dp;
        float s = extra;
#pragma unroll
        for (int i = 0; i < DMAX; i += 4) {
          if (i < dp) {
            float4 k4 = *reinterpret_cast<const float4*>(kp + i);
            s = fmaf(q[i], k4.x, s);
            s = fmaf(q[i + 1], k4.y, s);
            s = fmaf(q[i + 2], k4.z, s);
            s = fmaf(q[i + 3], k4.w, s);
          }
        }
        s += brow[kx];
        float p;
        if (s > mx) {               // rescale running sums to the new maximum
          float sc = expf(mx - s);  // exp(-inf) = 0 on the first key
          l *= sc;
#pragma unroll
          for (int i = 0; i < DMAX; i++) acc[i] *= sc;
          mx = s;
          p = 1.f;
        } else {
          p = expf(s - mx);
        }
        l += p;
        const float* vp = sv + (kr * ows + kx) * dp;
#pragma unroll
        for (int i = 0; i < DMAX; i += 4) {
          if (i < dp) {
            float4 v4 = *reinterpret_cast<const float4*>(vp + i);
            acc[i] = fmaf(p, v4.x, acc[i]);
            acc[i + 1] = fmaf(p, v4.y, acc[i + 1]);
            acc[i + 2] = fmaf(p, v4.z, acc[i + 2]);
            acc[i + 3] = fmaf(p, v4.w, acc[i + 3]);
          }
        }
      }
    }
  }
  if (!active) return;
  float inv = 1.f / l;
  T* orow = out + qtok * out_pitch;
#pragma unroll
  for (int i = 0; i < DMAX; i++)
    if (i < d) orow[head * d + i] = from_f<T>(acc[i] * inv);
  if (head == heads - 1)
    for (int c = C; c < out_zero_to; c++) orow[c] = from_f<T>(0.f);
}

template <typename T>
static int launch_t(const AttnArgs& a, cudaStream_t st) {
  int d = a.C / a.heads;
  int dp = (d + 3) & ~3;
  int Tn = a.ws + a.ows - 1;
  size_t smem = ((size_t)((Tn * Tn + 3) & ~3) + 2 * (size_t)KCHUNK_ROWS * a.ows * dp) * sizeof(float);
  dim3 grid(a.B * (a.H / a.ws) * (a.W / a.ws), a.heads);
  int threads = round_up(a.ws * a.ws, 32);
  const int pad = a.pad >= 0 ? a.pad : (a.ows - a.ws) / 2;
  if (d <= 32) {
    HB_CHECK_CUDA(cudaFuncSetAttribute(attn_simt_kernel<T, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attn_simt_kernel<T, 32><<<grid, threads, smem, st>>>((const T*)a.qkv, a.pitch, a.H, a.W, a.C, a.heads, a.ws, a.ows,
                                                         a.bias, (T*)a.out, a.out_pitch, a.out_zero_to, pad, a.kadd);
  } else {
    HB_CHECK_CUDA(cudaFuncSetAttribute(attn_simt_kernel<T, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attn_simt_kernel<T, 64><<<grid, threads, smem, st>>>((const T*)a.qkv, a.pitch, a.H, a.W, a.C, a.heads, a.ws, a.ows,
                                                         a.bias, (T*)a.out, a.out_pitch, a.out_zero_to, pad, a.kadd);
  }
  HB_CHECK_LAUNCH();
  return 0;
}

int launch_attn_simt(int dt, const AttnArgs& a, cudaStream_t st) {
  int d = a.C / a.heads;
  HB_REQUIRE(a.C % a.heads == 0 && d <= 64, HATB200_EUNSUPPORTED, "attention: head_dim=%d unsupported (<=64)", d);
  HB_REQUIRE(a.ws * a.ws <= 256, HATB200_EUNSUPPORTED, "attention: window_size=%d unsupported (<=16)", a.ws);
  HB_REQUIRE(a.H % a.ws == 0 && a.W % a.ws == 0 && (a.pad >= 0 || ((a.ows - a.ws) & 1) == 0), HATB200_EINVAL,
             "attention: bad window geometry");
  if (dt == HATB200_F32) return launch_t<float>(a, st);
  if (dt == HATB200_BF16) return launch_t<bf16>(a, st);
  set_error("attention: dtype %d unsupported", dt);
  return HATB200_EINVAL;
}

// ---------------------------------------------------------------------------
// compact[h][(dy+ws-1)*Tn + dx+ws-1] = table[wrap(rpi[q][k])][h]  with dy = ky-qy,
// dx = kx-qx.  "wrap" is PyTorch negative-index semantics (idx<0 -> idx+rows),
// hat_arch.py:378.  Every (q,k) pair writes its slot; a second pass checks that all
// pairs mapping to one slot agree (the index buffer is translation invariant for
// any table produced by calculate_rpi_oca, hat_arch.py:783-803).
// ---------------------------------------------------------------------------
template <typename T>
__global__ void pack_bias_kernel(const T* __restrict__ table, const int64_t* __restrict__ rpi, int heads, int ws,
                                 int ows, float* __restrict__ compact, int* __restrict__ err, int pass) {
  int nq = ws * ws, nk = ows * ows, Tn = ws + ows - 1, rows = Tn * Tn;
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nq * nk) return;
  int q = idx / nk, k = idx % nk;
  int qy = q / ws, qx = q % ws, ky = k / ows, kx = k % ows;
  long long r = rpi[idx];
  if (r < 0) r += rows;
  if (r < 0 || r >= rows) { atomicExch(err, 2); return; }
  int slot = (ky - qy + ws - 1) * Tn + (kx - qx + ws - 1);
  for (int h = 0; h < heads; h++) {
    float v = to_f(table[(size_t)r * heads + h]);
    if (pass == 0) compact[(size_t)h * rows + slot] = v;
    else if (compact[(size_t)h * rows + slot] != v) atomicExch(err, 1);
  }
}

int launch_pack_ocab_bias(const void* table, int table_dt, const int64_t* rpi, int heads, int ws, int ows,
                          float* compact, int* err_flag, cudaStream_t st) {
  int total = ws * ws * ows * ows;
  for (int pass = 0; pass < 2; pass++) {
    if (table_dt == HATB200_F32)
      pack_bias_kernel<float><<<cdiv(total, 256), 256, 0, st>>>((const float*)table, rpi, heads, ws, ows, compact,
                                                                err_flag, pass);
    else if (table_dt == HATB200_BF16)
      pack_bias_kernel<bf16><<<cdiv(total, 256), 256, 0, st>>>((const bf16*)table, rpi, heads, ws, ows, compact,
                                                               err_flag, pass);
    else {
      set_error("pack_ocab_bias: dtype %d unsupported", table_dt);
      return HATB200_EINVAL;
    }
    HB_CHECK_LAUNCH();
  }
  return 0;
}

}  // namespace hb
