// tapconv_simt.cu -- CUDA-core dense conv / Linear over channels-last tokens (fp32 math).
// This is the arithmetic of the fp32 parity mode (max-abs <= 1e-4 needs true fp32
// products; tcgen05 has no fp32 MMA) and the device-side cross-check for the
// tcgen05 kernel.  out[m, n] = sum_tap sum_c A[pix(m)+tap, c] * W[tap][c][n].
// Replaces F.conv2d / F.linear call sites hat_arch.py:84-86, :111, :117, :347-351,
// :391-392, :544, :854-856 and esc_arch.py:122-123, :144.
#include "kernels.cuh"

namespace hb {

template <typename T>
__device__ __forceinline__ void epilogue_store(const Epilogue& ep, int m, int n, int N, float v, int H, int W) {
  if (n >= N) {
    if (ep.out_t != nullptr && ep.shuffle_r == 1 && n < ep.out_t_zero_to)
      ((T*)ep.out_t)[(size_t)m * ep.out_t_pitch + ep.out_t_coff + n] = from_f<T>(0.f);
    return;
  }
  int hw = H * W;
  int b = m / hw;
  if (ep.bias) v += ep.bias[n];
  if (ep.dyn_k) {  // per-sample dynamic depthwise 3x3 on the compact first-pdim channels (esc_arch.py:123)
    int rem = m - b * hw;
    int py = rem / W, px = rem - py * W;
    const float* k = ep.dyn_k + ((size_t)b * N + n) * 9;
    const T* x = (const T*)ep.dyn_x;
    float a = 0.f;
#pragma unroll
    for (int ky = 0; ky < 3; ky++) {
      int yy = py + ky - 1;
      if (yy < 0 || yy >= H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; kx++) {
        int xx = px + kx - 1;
        if (xx < 0 || xx >= W) continue;
        a = fmaf(to_f(x[(size_t)(m + (ky - 1) * W + (kx - 1)) * ep.dyn_pitch + n]), k[ky * 3 + kx], a);
      }
    }
    v += a;
  }
  if (n < ep.scale_ncols) v *= ep.out_scale;
  if (ep.act == ACT_GELU) v = gelu_f(v);
  else if (ep.act == ACT_LRELU) v = lrelu_f(v);
  if (ep.eca_u)
    v += ep.eca_coef * ep.eca_scale[(size_t)b * N + n] * to_f(((const T*)ep.eca_u)[(size_t)m * ep.eca_pitch + n]);
  if (ep.resid) v += ep.resid[(size_t)m * ep.resid_pitch + n];
  if (ep.out_f32) ep.out_f32[(size_t)m * ep.out_f32_pitch + n] = v;
  if (ep.out_t) {
    if (ep.shuffle_r == 1) {
      ((T*)ep.out_t)[(size_t)m * ep.out_t_pitch + ep.out_t_coff + n] = from_f<T>(v);
    } else {  // PixelShuffle(r): out[b, c, y*r+i, x*r+j] = in[b, c*r*r + i*r + j, y, x]
      int r = ep.shuffle_r, rr = r * r;
      int c = n / rr, ij = n - c * rr, i = ij / r, j = ij - i * r;
      int rem = m - b * hw;
      int py = rem / W, px = rem - py * W;
      size_t ot = ((size_t)b * H * r + (size_t)py * r + i) * ((size_t)W * r) + (size_t)px * r + j;
      ((T*)ep.out_t)[ot * ep.out_t_pitch + ep.out_t_coff + c] = from_f<T>(v);
    }
  }
}

constexpr int SBK = 16;

template <typename T, int BM, int BN, int TM, int TN>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
tapconv_simt_kernel(const T* __restrict__ A, int a_pitch, int a_coff, int Cin, int Bn, int H, int W, int KH, int KW,
                    const float* __restrict__ Wt, int Np, int N, Epilogue ep) {
  constexpr int NT = (BM / TM) * (BN / TN);
  __shared__ float As[SBK][BM + 4];
  __shared__ float Bs[SBK][BN + 4];
  const int M = Bn * H * W;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; i++)
#pragma unroll
    for (int j = 0; j < TN; j++) acc[i][j] = 0.f;

  // each thread stages A elements (row = e / SBK, k = e % SBK), e = tid, tid + NT, ...
  constexpr int A_PER = (BM * SBK + NT - 1) / NT;
  int a_row[A_PER], a_py[A_PER], a_px[A_PER];
#pragma unroll
  for (int i = 0; i < A_PER; i++) {
    int e = tid + i * NT;
    int row = e / SBK;
    int m = m0 + row;
    a_row[i] = (e < BM * SBK && m < M) ? m : -1;
    int rem = m % (H * W);
    a_py[i] = rem / W;
    a_px[i] = rem % W;
  }
  const int ph = KH / 2, pw = KW / 2;
  for (int tap = 0; tap < KH * KW; tap++) {
    const int dy = tap / KW - ph, dx = tap % KW - pw;
    const float* wt = Wt + (size_t)tap * Cin * Np;
    for (int k0 = 0; k0 < Cin; k0 += SBK) {
#pragma unroll
      for (int i = 0; i < A_PER; i++) {
        int e = tid + i * NT;
        if (e < BM * SBK) {
          int row = e / SBK, kk = e % SBK;
          float v = 0.f;
          int yy = a_py[i] + dy, xx = a_px[i] + dx;
          if (a_row[i] >= 0 && k0 + kk < Cin && yy >= 0 && yy < H && xx >= 0 && xx < W)
            v = to_f(A[(size_t)(a_row[i] + dy * W + dx) * a_pitch + a_coff + k0 + kk]);
          As[kk][row] = v;
        }
      }
      for (int e = tid; e < SBK * BN; e += NT) {
        int kk = e / BN, nn = e % BN;
        Bs[kk][nn] = (k0 + kk < Cin && n0 + nn < Np) ? wt[(size_t)(k0 + kk) * Np + n0 + nn] : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < SBK; kk++) {
        float a[TM], b[TN];
#pragma unroll
        for (int i = 0; i < TM; i++) a[i] = As[kk][ty * TM + i];
#pragma unroll
        for (int j = 0; j < TN; j++) b[j] = Bs[kk][tx * TN + j];
#pragma unroll
        for (int i = 0; i < TM; i++)
#pragma unroll
          for (int j = 0; j < TN; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < TM; i++) {
    int m = m0 + ty * TM + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < TN; j++) epilogue_store<T>(ep, m, n0 + tx * TN + j, N, acc[i][j], H, W);
  }
}

template <typename T>
static int launch_t(const ConvArgs& a, const float* w, int Np, const Epilogue& ep, cudaStream_t st) {
  int M = a.B * a.H * a.W;
  int ncols = ep.out_t_zero_to > a.N ? ep.out_t_zero_to : a.N;
  if (a.N <= 16 && ncols <= 16) {
    dim3 grid(cdiv(M, 64), 1);
    tapconv_simt_kernel<T, 64, 16, 1, 4><<<grid, 256, 0, st>>>((const T*)a.A, a.a_pitch, a.a_coff, a.Cin, a.B, a.H,
                                                               a.W, a.KH, a.KW, w, Np, a.N, ep);
  } else {
    dim3 grid(cdiv(M, 64), cdiv(ncols, 64));
    tapconv_simt_kernel<T, 64, 64, 4, 4><<<grid, 256, 0, st>>>((const T*)a.A, a.a_pitch, a.a_coff, a.Cin, a.B, a.H,
                                                               a.W, a.KH, a.KW, w, Np, a.N, ep);
  }
  HB_CHECK_LAUNCH();
  return 0;
}

int launch_tapconv_simt(int dt, const ConvArgs& a, const float* w, int Np, const Epilogue& ep, cudaStream_t st) {
  HB_REQUIRE(ep.ln_out == nullptr && ep.colsum == nullptr, HATB200_EINVAL,
             "tapconv_simt: fused LayerNorm / column sums are tensor-core epilogue features");
  HB_REQUIRE((a.KH & 1) && (a.KW & 1), HATB200_EINVAL, "tapconv: even kernel %dx%d", a.KH, a.KW);
  if (dt == HATB200_F32) return launch_t<float>(a, w, Np, ep, st);
  if (dt == HATB200_BF16) return launch_t<bf16>(a, w, Np, ep, st);
  set_error("tapconv_simt: dtype %d unsupported", dt);
  return HATB200_EINVAL;
}

}  // namespace hb
