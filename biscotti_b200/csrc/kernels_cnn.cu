// K2: batched-over-peers local step of the reference's one-convolution CNNs.
//
//   mnist_cnn : Conv2d(1,16,5,1,pad 4) + ReLU + MaxPool(1) + Linear(16*32*32, 10)   ML/Pytorch/mnist_cnn_model.py:12-39
//   cifar_cnn : Conv2d(3,20,3,1,pad 3) + ReLU + MaxPool(1) + Linear(25920, 10)      ML/Pytorch/cifar_cnn_model.py:11-29
//   step      : one minibatch, CrossEntropyLoss, clip_grad_norm 100, flatten in named_parameters order
//               (conv.weight, conv.bias, fc.weight, fc.bias), update = -grad    ML/Pytorch/client.py:38-65
//
// Every peer starts from the same global weights, so the forward convolution is one launch over all
// P*B images; the backward pass is per-peer (each peer needs its own weight gradient).  Five kernels:
//   cnn1_conv_fwd     activations act[p][b][oc][oy][ox] = relu(conv)
//   cnn1_fc_fwd       logits (one CTA per image)
//   cnn1_loss         softmax, dlogits, loss, fc.bias gradient (one warp per peer)
//   cnn1_fc_bwd       fc.weight gradient (the bulk of the update: C*J floats per peer), d(act) masked by
//                     ReLU written over act, partial squared norms
//   cnn1_conv_wgrad   conv.weight / conv.bias gradients (one CTA per (peer, out-channel)), partial norms
//   cnn1_coef/scale   fixed-shape reduction of the partial norms -> clip coefficient, U = -coef * grad
// Partial norms are written to per-CTA slots and summed in index order: the clip coefficient (and with it
// the ledger) is bit-reproducible.
#include <string.h>
#include "common.cuh"

namespace bsc {

struct Cnn1Shape {
  int IC, OC, K, pad, H, W;   // input channels, output channels, kernel, padding, input height/width
  int OH, OW, J, C, B;        // output height/width, J = OC*OH*OW, classes, batch
  int off_cw, off_cb, off_fw, off_fb, d;   // offsets of the four parameter tensors in the flat vector
};

__global__ void __launch_bounds__(256) cnn1_conv_fwd_kernel(const float* __restrict__ w, const float* __restrict__ X,
                                                            float* __restrict__ act, Cnn1Shape s, int PB) {
  extern __shared__ float sw[];   // conv weights + bias
  const int nw = s.OC * s.IC * s.K * s.K;
  for (int i = threadIdx.x; i < nw + s.OC; i += blockDim.x) sw[i] = i < nw ? w[s.off_cw + i] : w[s.off_cb + i - nw];
  __syncthreads();
  const long long total = (long long)PB * s.J;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int ox = (int)(t % s.OW);
    const int oy = (int)((t / s.OW) % s.OH);
    const int oc = (int)((t / ((long long)s.OW * s.OH)) % s.OC);
    const long long img = t / s.J;
    const float* x = X + img * s.IC * s.H * s.W;
    float acc = sw[nw + oc];
    for (int ic = 0; ic < s.IC; ic++)
      for (int ky = 0; ky < s.K; ky++) {
        const int iy = oy + ky - s.pad;
        if (iy < 0 || iy >= s.H) continue;
        for (int kx = 0; kx < s.K; kx++) {
          const int ix = ox + kx - s.pad;
          if (ix < 0 || ix >= s.W) continue;
          acc = fmaf(sw[((oc * s.IC + ic) * s.K + ky) * s.K + kx], x[(ic * s.H + iy) * s.W + ix], acc);
        }
      }
    act[t] = acc > 0.f ? acc : 0.f;
  }
}

// grid (B, P): the C logits of one image (one warp per class, strided), written to logits[p][b][c]
__global__ void __launch_bounds__(256) cnn1_fc_fwd_kernel(const float* __restrict__ w, const float* __restrict__ act,
                                                          float* __restrict__ logits, Cnn1Shape s) {
  const int b = blockIdx.x, p = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const float* ar = act + ((size_t)p * s.B + b) * s.J;
  for (int c = warp; c < s.C; c += nwarp) {
    const float* wr = w + s.off_fw + (size_t)c * s.J;
    float acc = 0.f;
    for (int j = lane; j < s.J; j += 32) acc = fmaf(ar[j], wr[j], acc);
    acc = warp_sum(acc);
    if (lane == 0) logits[((size_t)p * s.B + b) * s.C + c] = acc + w[s.off_fb + c];
  }
}
// one warp per peer: softmax, dl = (softmax - onehot)/B (written over logits), loss, fc.bias gradient
__global__ void cnn1_loss_kernel(float* __restrict__ dl, const int* __restrict__ y, float* __restrict__ loss,
                                 float* __restrict__ U, int ldu, int P, Cnn1Shape s) {
  const int p = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (p >= P) return;
  float* sl = dl + (size_t)p * s.B * s.C;
  float myloss = 0.f;
  if (lane < s.B) {
    float m = -INFINITY;
    for (int c = 0; c < s.C; c++) m = fmaxf(m, sl[lane * s.C + c]);
    float sum = 0.f;
    for (int c = 0; c < s.C; c++) sum += expf(sl[lane * s.C + c] - m);
    const float lse = m + logf(sum);
    const int lab = y[p * s.B + lane];
    myloss = (lse - sl[lane * s.C + lab]) / (float)s.B;
    for (int c = 0; c < s.C; c++) {
      const float pr = expf(sl[lane * s.C + c] - lse);
      sl[lane * s.C + c] = (pr - (c == lab ? 1.f : 0.f)) / (float)s.B;
    }
  }
  myloss = warp_sum(myloss);
  if (lane == 0 && loss) loss[p] = myloss;
  __syncwarp();
  if (lane < s.C) {
    float g = 0.f;
    for (int b = 0; b < s.B; b++) g += sl[b * s.C + lane];
    U[(size_t)p * ldu + s.off_fb + lane] = g;   // unscaled gradient; cnn1_scale finishes it
  }
}

// grid (ceil(J/256), P): thread j -> fc.weight gradient column j, masked d(act) written over act
__global__ void __launch_bounds__(256) cnn1_fc_bwd_kernel(const float* __restrict__ w, float* __restrict__ act,
                                                          const float* __restrict__ dl, float* __restrict__ U, int ldu,
                                                          float* __restrict__ partial, int nslots, Cnn1Shape s) {
  __shared__ float sdl[32 * 16];
  __shared__ float red[40];
  const int p = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
  for (int i = threadIdx.x; i < s.B * s.C; i += blockDim.x) sdl[i] = dl[(size_t)p * s.B * s.C + i];
  __syncthreads();
  float nrm = 0.f;
  if (j < s.J) {
    float wc[16];
    for (int c = 0; c < s.C; c++) wc[c] = w[s.off_fw + (size_t)c * s.J + j];
    float g[16];
    for (int c = 0; c < s.C; c++) g[c] = 0.f;
    float* a = act + (size_t)p * s.B * s.J + j;
    for (int b = 0; b < s.B; b++) {
      const float av = a[(size_t)b * s.J];
      float da = 0.f;
      for (int c = 0; c < s.C; c++) {
        const float d = sdl[b * s.C + c];
        g[c] = fmaf(d, av, g[c]);
        da = fmaf(d, wc[c], da);
      }
      a[(size_t)b * s.J] = av > 0.f ? da : 0.f;   // d(pre-activation)
    }
    for (int c = 0; c < s.C; c++) {
      U[(size_t)p * ldu + s.off_fw + (size_t)c * s.J + j] = g[c];
      nrm = fmaf(g[c], g[c], nrm);
    }
  }
  const float t = block_sum(nrm, red);
  if (threadIdx.x == 0) partial[(size_t)p * nslots + blockIdx.x] = t;
}

// grid (OC, P): conv.weight[oc][*] and conv.bias[oc] gradients of one peer
__global__ void __launch_bounds__(256) cnn1_conv_wgrad_kernel(const float* __restrict__ X, const float* __restrict__ dpre,
                                                              float* __restrict__ U, int ldu,
                                                              float* __restrict__ partial, int nslots, int slot0,
                                                              Cnn1Shape s) {
  __shared__ float red[40];
  const int oc = blockIdx.x, p = blockIdx.y, tid = threadIdx.x;
  const int taps = s.IC * s.K * s.K;   // <= 75 for the reference models
  float acc[80];
  for (int t = 0; t < taps; t++) acc[t] = 0.f;
  float accb = 0.f;
  const int npos = s.B * s.OH * s.OW;
  for (int pos = tid; pos < npos; pos += blockDim.x) {
    const int ox = pos % s.OW, oy = (pos / s.OW) % s.OH, b = pos / (s.OW * s.OH);
    const float d = dpre[((size_t)(p * s.B + b) * s.OC + oc) * s.OH * s.OW + oy * s.OW + ox];
    if (d == 0.f) continue;
    accb += d;
    const float* x = X + (size_t)(p * s.B + b) * s.IC * s.H * s.W;
    int t = 0;
    for (int ic = 0; ic < s.IC; ic++)
      for (int ky = 0; ky < s.K; ky++) {
        const int iy = oy + ky - s.pad;
        for (int kx = 0; kx < s.K; kx++, t++) {
          const int ix = ox + kx - s.pad;
          if (iy >= 0 && iy < s.H && ix >= 0 && ix < s.W) acc[t] = fmaf(d, x[(ic * s.H + iy) * s.W + ix], acc[t]);
        }
      }
  }
  float nrm = 0.f;
  for (int t = 0; t < taps; t++) {
    const float g = block_sum(acc[t], red);
    if (tid == 0) {
      U[(size_t)p * ldu + s.off_cw + oc * taps + t] = g;
      nrm = fmaf(g, g, nrm);
    }
  }
  const float gb = block_sum(accb, red);
  if (tid == 0) {
    U[(size_t)p * ldu + s.off_cb + oc] = gb;
    nrm = fmaf(gb, gb, nrm);
    partial[(size_t)p * nslots + slot0 + oc] = nrm;
  }
}

// one CTA per peer: clip coefficient from the partial squared norms (+ fc.bias), fixed-shape tree reduction
// (deterministic: the same slots are combined in the same order on every run and every rank)
__global__ void __launch_bounds__(256) cnn1_coef_kernel(const float* __restrict__ U, int ldu, const float* __restrict__ partial,
                                                        int nslots, float clip, float* __restrict__ coef, Cnn1Shape s) {
  __shared__ float red[40];
  const int p = blockIdx.x;
  float t = 0.f;
  for (int i = threadIdx.x; i < nslots; i += blockDim.x) t += partial[(size_t)p * nslots + i];
  if (threadIdx.x < s.C) {
    const float g = U[(size_t)p * ldu + s.off_fb + threadIdx.x];
    t = fmaf(g, g, t);
  }
  const float tot = block_sum(t, red);
  if (threadIdx.x == 0) {
    const float c = clip / (sqrtf(tot) + 1e-6f);
    coef[p] = c > 1.f ? 1.f : c;
  }
}
// grid (ceil(ldu/4096), P): U = -coef * grad, padding zeroed
__global__ void __launch_bounds__(256) cnn1_scale_kernel(float* __restrict__ U, int ldu, const float* __restrict__ coef,
                                                         Cnn1Shape s) {
  const int p = blockIdx.y;
  const float m = -coef[p];
  float* row = U + (size_t)p * ldu;
  const int k0 = blockIdx.x * 4096;
  for (int k = k0 + threadIdx.x; k < min(ldu, k0 + 4096); k += blockDim.x) row[k] = k < s.d ? row[k] * m : 0.f;
}

// ---------------------------------------------------------------------------------------------------
// Two-convolution model with 2x2 max-pooling after each ReLU:
//   lfw_cnn : Conv2d(3,18,3,1,1)+ReLU+MaxPool2 + Conv2d(18,36,3,1,1)+ReLU+MaxPool2 + Linear(5940, 2)
//             ML/Pytorch/lfw_cnn_model.py:11-29 (18 254 parameters)
// The pooled activation and a 2-bit arg-max are all the forward pass keeps per 2x2 window: relu(max) ==
// max(relu), and the backward pass routes a gradient to the arg-max position only where the pooled value is
// positive.  The classifier head reuses the cnn1_fc_* / loss / coef / scale kernels above (J = pooled size).

// thread -> one pooled output (img, oc, py, px): the four conv outputs of its window from one (K+1)^2 patch per channel
template <int K>
__global__ void __launch_bounds__(256) cnn2_conv_pool_fwd_kernel(const float* __restrict__ wt, const float* __restrict__ bias,
                                                                 const float* __restrict__ in, float* __restrict__ out,
                                                                 uint8_t* __restrict__ arg, int IC, int H, int W, int OC,
                                                                 int pad, int PH, int PW, long long total) {
  extern __shared__ float sw[];
  const int nw = OC * IC * K * K;
  for (int i = threadIdx.x; i < nw + OC; i += blockDim.x) sw[i] = i < nw ? wt[i] : bias[i - nw];
  __syncthreads();
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int px = (int)(t % PW);
    const int py = (int)((t / PW) % PH);
    const int oc = (int)((t / ((long long)PW * PH)) % OC);
    const long long img = t / ((long long)PW * PH * OC);
    const float* x = in + img * IC * H * W;
    const float b = sw[nw + oc];
    float a00 = b, a01 = b, a10 = b, a11 = b;
    const int iy0 = 2 * py - pad, ix0 = 2 * px - pad;
    for (int ic = 0; ic < IC; ic++) {
      float patch[K + 1][K + 1];
#pragma unroll
      for (int r = 0; r <= K; r++)
#pragma unroll
        for (int c = 0; c <= K; c++) {
          const int iy = iy0 + r, ix = ix0 + c;
          patch[r][c] = (iy >= 0 && iy < H && ix >= 0 && ix < W) ? x[(ic * H + iy) * W + ix] : 0.f;
        }
      const float* wk = sw + (oc * IC + ic) * K * K;
#pragma unroll
      for (int ky = 0; ky < K; ky++)
#pragma unroll
        for (int kx = 0; kx < K; kx++) {
          const float wv = wk[ky * K + kx];
          a00 = fmaf(wv, patch[ky][kx], a00);
          a01 = fmaf(wv, patch[ky][kx + 1], a01);
          a10 = fmaf(wv, patch[ky + 1][kx], a10);
          a11 = fmaf(wv, patch[ky + 1][kx + 1], a11);
        }
    }
    float best = a00;
    int q = 0;
    if (a01 > best) { best = a01; q = 1; }
    if (a10 > best) { best = a10; q = 2; }
    if (a11 > best) { best = a11; q = 3; }
    out[t] = best > 0.f ? best : 0.f;
    arg[t] = (uint8_t)q;
  }
}

// grid (OC, P): weight / bias gradients of one output channel of one peer from the pooled gradient dp (already
// masked by the ReLU) and the arg-max map; input channels are processed ICB at a time to bound registers.
template <int K, int ICB>
__global__ void __launch_bounds__(256) cnn2_conv_wgrad_kernel(const float* __restrict__ in, const float* __restrict__ dp,
                                                              const uint8_t* __restrict__ arg, float* __restrict__ U, int ldu,
                                                              int off_w, int off_b, float* __restrict__ partial, int nslots,
                                                              int slot0, int B, int IC, int H, int W, int OC, int pad, int PH,
                                                              int PW) {
  __shared__ float red[40];
  const int oc = blockIdx.x, p = blockIdx.y, tid = threadIdx.x;
  const int npos = B * PH * PW;
  float nrm = 0.f;
  for (int ic0 = 0; ic0 < IC; ic0 += ICB) {
    float acc[ICB * K * K];
#pragma unroll
    for (int t = 0; t < ICB * K * K; t++) acc[t] = 0.f;
    float accb = 0.f;
    for (int pos = tid; pos < npos; pos += blockDim.x) {
      const int px = pos % PW, py = (pos / PW) % PH, b = pos / (PW * PH);
      const size_t idx = ((size_t)(p * B + b) * OC + oc) * PH * PW + py * PW + px;
      const float d = dp[idx];
      if (d == 0.f) continue;
      const int a = arg[idx];
      const int oy = 2 * py + (a >> 1), ox = 2 * px + (a & 1);
      accb += d;
      const float* x = in + (size_t)(p * B + b) * IC * H * W;
#pragma unroll
      for (int i = 0; i < ICB; i++) {
        const int ic = ic0 + i;
        if (ic >= IC) break;
#pragma unroll
        for (int ky = 0; ky < K; ky++) {
          const int iy = oy + ky - pad;
#pragma unroll
          for (int kx = 0; kx < K; kx++) {
            const int ix = ox + kx - pad;
            if (iy >= 0 && iy < H && ix >= 0 && ix < W)
              acc[(i * K + ky) * K + kx] = fmaf(d, x[(ic * H + iy) * W + ix], acc[(i * K + ky) * K + kx]);
          }
        }
      }
    }
#pragma unroll
    for (int t = 0; t < ICB * K * K; t++) {
      const float g = block_sum(acc[t], red);
      if (tid == 0 && ic0 + t / (K * K) < IC) {
        U[(size_t)p * ldu + off_w + (oc * IC + ic0) * K * K + t] = g;
        nrm = fmaf(g, g, nrm);
      }
    }
    if (ic0 == 0) {
      const float gb = block_sum(accb, red);
      if (tid == 0) {
        U[(size_t)p * ldu + off_b + oc] = gb;
        nrm = fmaf(gb, gb, nrm);
      }
    }
  }
  if (tid == 0) partial[(size_t)p * nslots + slot0 + oc] = nrm;
}

// Dense form of the second convolution's output gradient: dense[img][oc][oy][ox] = dp2 at the arg-max position of
// each pooled cell, zero elsewhere (the caller zero-fills `dense`).  With it the data gradient below is a plain,
// branch-free correlation instead of a per-tap arg-max test.
__global__ void __launch_bounds__(256) cnn2_unpool_kernel(const float* __restrict__ dp2, const uint8_t* __restrict__ arg2,
                                                          float* __restrict__ dense, int PH, int PW, int OH, int OW,
                                                          long long total) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const float d = dp2[t];
    if (d == 0.f) continue;
    const int px = (int)(t % PW), py = (int)((t / PW) % PH);
    const long long plane = t / ((long long)PW * PH);            // img * OC + oc
    const int a = arg2[t];
    dense[plane * OH * OW + (size_t)(2 * py + (a >> 1)) * OW + (2 * px + (a & 1))] = d;
  }
}
// thread -> one input position (img, ic, y, x): g = sum_oc sum_taps dense[oc][y-ky+pad][x-kx+pad] * w[oc][ic][ky][kx],
// masked by the first layer's ReLU (a1 > 0)
template <int K>
__global__ void __launch_bounds__(256) cnn2_dgrad_dense_kernel(const float* __restrict__ wt, const float* __restrict__ dense,
                                                               const float* __restrict__ a1, float* __restrict__ dp1, int IC,
                                                               int H, int W, int OC, int pad, int OH, int OW, long long total) {
  extern __shared__ float sw[];
  const int nw = OC * IC * K * K;
  for (int i = threadIdx.x; i < nw; i += blockDim.x) sw[i] = wt[i];
  __syncthreads();
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    if (!(a1[t] > 0.f)) {
      dp1[t] = 0.f;
      continue;
    }
    const int x = (int)(t % W);
    const int y = (int)((t / W) % H);
    const int ic = (int)((t / ((long long)W * H)) % IC);
    const long long img = t / ((long long)W * H * IC);
    float g = 0.f;
    for (int oc = 0; oc < OC; oc++) {
      const float* dd = dense + ((size_t)img * OC + oc) * OH * OW;
      const float* wk = sw + (oc * IC + ic) * K * K;
#pragma unroll
      for (int ky = 0; ky < K; ky++) {
        const int oy = y - ky + pad;
        if (oy < 0 || oy >= OH) continue;
#pragma unroll
        for (int kx = 0; kx < K; kx++) {
          const int ox = x - kx + pad;
          if (ox >= 0 && ox < OW) g = fmaf(dd[oy * OW + ox], wk[ky * K + kx], g);
        }
      }
    }
    dp1[t] = g;
  }
}

}  // namespace bsc

using namespace bsc;

extern "C" {

// scratch: act [P*B*J] floats, dl [P*B*C], partial [P*nslots] with nslots = ceil(J/256) + OC, coef [P]
long long bsc_cnn1_scratch_floats(int P, int B, int IC, int OC, int K, int pad, int H, int W, int C) {
  const int OH = H + 2 * pad - K + 1, OW = W + 2 * pad - K + 1;
  const long long J = (long long)OC * OH * OW;
  const long long nslots = (J + 255) / 256 + OC;
  return (long long)P * B * J + (long long)P * B * C + (long long)P * nslots + P;
}

// One-conv-layer CNN local step for P peers.  w: flat global weights; X: [P][B][IC][H][W]; y: [P][B];
// U: [P][ldu] out (= -clip(grad)); loss: [P] (may be null); scratch: see bsc_cnn1_scratch_floats.
int bsc_cnn1_local_step(const float* w, const float* X, const int* y, float* U, float* loss, float* scratch, int P, int B,
                        int IC, int OC, int K, int pad, int H, int W, int C, int ldu, float clip, void* stream) {
  Cnn1Shape s;
  s.IC = IC; s.OC = OC; s.K = K; s.pad = pad; s.H = H; s.W = W;
  s.OH = H + 2 * pad - K + 1;
  s.OW = W + 2 * pad - K + 1;
  s.J = OC * s.OH * s.OW;
  s.C = C;
  s.B = B;
  s.off_cw = 0;
  s.off_cb = OC * IC * K * K;
  s.off_fw = s.off_cb + OC;
  s.off_fb = s.off_fw + C * s.J;
  s.d = s.off_fb + C;
  if (C > 16 || B > 32 || IC * K * K > 80 || ldu < s.d) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const int nfc = (s.J + 255) / 256;
  const int nslots = nfc + OC;
  float* act = scratch;
  float* dl = act + (size_t)P * B * s.J;
  float* partial = dl + (size_t)P * B * C;
  float* coef = partial + (size_t)P * nslots;
  const long long total = (long long)P * B * s.J;
  const int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  const size_t smem = (size_t)(OC * IC * K * K + OC) * sizeof(float);
  cnn1_conv_fwd_kernel<<<grid, 256, smem, st>>>(w, X, act, s, P * B);
  cnn1_fc_fwd_kernel<<<dim3(B, P), 256, 0, st>>>(w, act, dl, s);
  cnn1_loss_kernel<<<(P + 3) / 4, 128, 0, st>>>(dl, y, loss, U, ldu, P, s);
  cnn1_fc_bwd_kernel<<<dim3(nfc, P), 256, 0, st>>>(w, act, dl, U, ldu, partial, nslots, s);
  cnn1_conv_wgrad_kernel<<<dim3(OC, P), 256, 0, st>>>(X, act, U, ldu, partial, nslots, nfc, s);
  cnn1_coef_kernel<<<P, 256, 0, st>>>(U, ldu, partial, nslots, clip, coef, s);
  cnn1_scale_kernel<<<dim3((ldu + 4095) / 4096, P), 256, 0, st>>>(U, ldu, coef, s);
  return finish_launch("cnn1_local_step", 7);
}

// ---- two-convolution model (conv+ReLU+MaxPool2 twice, then a linear head) ------------------------------------
struct Cnn2Dims {
  int OH1, OW1, PH1, PW1, OH2, OW2, PH2, PW2;
  long long n1, J, nfc, nslots;   // n1 = OC1*PH1*PW1 floats per image after layer 1, J = OC2*PH2*PW2
};
static Cnn2Dims cnn2_dims(int H, int W, int OC1, int K1, int pad1, int OC2, int K2, int pad2) {
  Cnn2Dims m;
  m.OH1 = H + 2 * pad1 - K1 + 1; m.OW1 = W + 2 * pad1 - K1 + 1;
  m.PH1 = m.OH1 / 2; m.PW1 = m.OW1 / 2;
  m.OH2 = m.PH1 + 2 * pad2 - K2 + 1; m.OW2 = m.PW1 + 2 * pad2 - K2 + 1;
  m.PH2 = m.OH2 / 2; m.PW2 = m.OW2 / 2;
  m.n1 = (long long)OC1 * m.PH1 * m.PW1;
  m.J = (long long)OC2 * m.PH2 * m.PW2;
  m.nfc = (m.J + 255) / 256;
  m.nslots = m.nfc + OC2 + OC1;
  return m;
}
// scratch (floats): pool1 [PB*n1], dp1 [PB*n1], act [PB*J], dl [PB*C], partial [P*nslots], coef [P],
// arg1 bytes [PB*n1], arg2 bytes [PB*J] (each rounded up to whole floats)
long long bsc_cnn2_scratch_floats(int P, int B, int IC, int H, int W, int OC1, int K1, int pad1, int OC2, int K2, int pad2,
                                  int C) {
  const Cnn2Dims m = cnn2_dims(H, W, OC1, K1, pad1, OC2, K2, pad2);
  const long long PB = (long long)P * B;
  return 2 * PB * m.n1 + PB * m.J + PB * C + (long long)P * m.nslots + P + (PB * m.n1 + 3) / 4 + (PB * m.J + 3) / 4 + 8 +
         PB * OC2 * m.OH2 * m.OW2;   // + dense gradient of the second convolution's output
}

extern "C++" {
template <int K>
static void cnn2_launch_fwd(const float* wt, const float* bias, const float* in, float* out, uint8_t* arg, int IC, int H,
                            int W, int OC, int pad, int PH, int PW, long long total, cudaStream_t st) {
  const long long blocks = (total + 255) / 256;
  const int grid = (int)(blocks < 148 * 16 ? blocks : 148 * 16);
  const size_t smem = (size_t)(OC * IC * K * K + OC) * sizeof(float);
  cnn2_conv_pool_fwd_kernel<K><<<grid, 256, smem, st>>>(wt, bias, in, out, arg, IC, H, W, OC, pad, PH, PW, total);
}
template <int K>
static void cnn2_launch_wgrad(const float* in, const float* dp, const uint8_t* arg, float* U, int ldu, int off_w, int off_b,
                              float* partial, int nslots, int slot0, int P, int B, int IC, int H, int W, int OC, int pad,
                              int PH, int PW, cudaStream_t st) {
  cnn2_conv_wgrad_kernel<K, 3><<<dim3(OC, P), 256, 0, st>>>(in, dp, arg, U, ldu, off_w, off_b, partial, nslots, slot0, B, IC,
                                                            H, W, OC, pad, PH, PW);
}
}  // extern "C++"
#define CNN2_BY_K(K, call3, call5) \
  do {                             \
    if ((K) == 3) { call3; } else { call5; } \
  } while (0)

// Local step of the two-convolution model for P peers.  w: flat global weights in named_parameters order
// (conv0.weight, conv0.bias, conv1.weight, conv1.bias, fc.weight, fc.bias); X: [P][B][IC][H][W]; y: [P][B];
// U: [P][ldu] out (= -clip(grad)); loss: [P] or null; scratch: bsc_cnn2_scratch_floats floats.
int bsc_cnn2_local_step(const float* w, const float* X, const int* y, float* U, float* loss, float* scratch, int P, int B,
                        int IC, int H, int W, int OC1, int K1, int pad1, int OC2, int K2, int pad2, int C, int ldu,
                        float clip, void* stream) {
  if ((K1 != 3 && K1 != 5) || (K2 != 3 && K2 != 5) || C > 16 || B > 32) return -1;
  const Cnn2Dims m = cnn2_dims(H, W, OC1, K1, pad1, OC2, K2, pad2);
  if (m.PH2 <= 0 || m.PW2 <= 0) return -1;
  const int off_w1 = 0, off_b1 = OC1 * IC * K1 * K1, off_w2 = off_b1 + OC1, off_b2 = off_w2 + OC2 * OC1 * K2 * K2;
  Cnn1Shape s;   // the head kernels only read J, C, B, off_fw, off_fb, d
  memset(&s, 0, sizeof(s));
  s.J = (int)m.J; s.C = C; s.B = B;
  s.off_fw = off_b2 + OC2;
  s.off_fb = s.off_fw + C * (int)m.J;
  s.d = s.off_fb + C;
  if (ldu < s.d || (size_t)(OC2 * OC1 * K2 * K2 + OC2) * sizeof(float) > 48 * 1024) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const long long PB = (long long)P * B;
  const int nslots = (int)m.nslots, nfc = (int)m.nfc;
  float* pool1 = scratch;
  float* dp1 = pool1 + PB * m.n1;
  float* act = dp1 + PB * m.n1;
  float* dl = act + PB * m.J;
  float* partial = dl + PB * C;
  float* coef = partial + (long long)P * nslots;
  uint8_t* arg1 = (uint8_t*)(coef + P);
  uint8_t* arg2 = arg1 + ((PB * m.n1 + 3) / 4) * 4;
  float* dense = (float*)(arg2 + ((PB * m.J + 3) / 4) * 4 + 16);   // [PB][OC2][OH2][OW2]
  // forward
  CNN2_BY_K(K1, cnn2_launch_fwd<3>(w + off_w1, w + off_b1, X, pool1, arg1, IC, H, W, OC1, pad1, m.PH1, m.PW1, PB * m.n1, st),
            cnn2_launch_fwd<5>(w + off_w1, w + off_b1, X, pool1, arg1, IC, H, W, OC1, pad1, m.PH1, m.PW1, PB * m.n1, st));
  CNN2_BY_K(K2, cnn2_launch_fwd<3>(w + off_w2, w + off_b2, pool1, act, arg2, OC1, m.PH1, m.PW1, OC2, pad2, m.PH2, m.PW2, PB * m.J, st),
            cnn2_launch_fwd<5>(w + off_w2, w + off_b2, pool1, act, arg2, OC1, m.PH1, m.PW1, OC2, pad2, m.PH2, m.PW2, PB * m.J, st));
  cnn1_fc_fwd_kernel<<<dim3(B, P), 256, 0, st>>>(w, act, dl, s);
  cnn1_loss_kernel<<<(P + 3) / 4, 128, 0, st>>>(dl, y, loss, U, ldu, P, s);
  // backward: head (act <- masked gradient of the pooled layer-2 output), then the two convolutions
  cnn1_fc_bwd_kernel<<<dim3(nfc, P), 256, 0, st>>>(w, act, dl, U, ldu, partial, nslots, s);
  CNN2_BY_K(K2, cnn2_launch_wgrad<3>(pool1, act, arg2, U, ldu, off_w2, off_b2, partial, nslots, nfc, P, B, OC1, m.PH1, m.PW1, OC2, pad2, m.PH2, m.PW2, st),
            cnn2_launch_wgrad<5>(pool1, act, arg2, U, ldu, off_w2, off_b2, partial, nslots, nfc, P, B, OC1, m.PH1, m.PW1, OC2, pad2, m.PH2, m.PW2, st));
  {
    // data gradient of the second convolution: expand the pooled gradient to its dense form, then a plain correlation
    const long long nd = PB * OC2 * (long long)m.OH2 * m.OW2;
    cudaMemsetAsync(dense, 0, (size_t)nd * sizeof(float), st);
    const long long nj = PB * m.J;
    const int g1 = (int)((nj + 255) / 256 < 148 * 16 ? (nj + 255) / 256 : 148 * 16);
    cnn2_unpool_kernel<<<g1, 256, 0, st>>>(act, arg2, dense, m.PH2, m.PW2, m.OH2, m.OW2, nj);
    const long long n1 = PB * m.n1;
    const int g2 = (int)((n1 + 255) / 256 < 148 * 16 ? (n1 + 255) / 256 : 148 * 16);
    const size_t smem = (size_t)OC2 * OC1 * K2 * K2 * sizeof(float);
    if (K2 == 3)
      cnn2_dgrad_dense_kernel<3><<<g2, 256, smem, st>>>(w + off_w2, dense, pool1, dp1, OC1, m.PH1, m.PW1, OC2, pad2, m.OH2, m.OW2, n1);
    else
      cnn2_dgrad_dense_kernel<5><<<g2, 256, smem, st>>>(w + off_w2, dense, pool1, dp1, OC1, m.PH1, m.PW1, OC2, pad2, m.OH2, m.OW2, n1);
  }
  CNN2_BY_K(K1, cnn2_launch_wgrad<3>(X, dp1, arg1, U, ldu, off_w1, off_b1, partial, nslots, nfc + OC2, P, B, IC, H, W, OC1, pad1, m.PH1, m.PW1, st),
            cnn2_launch_wgrad<5>(X, dp1, arg1, U, ldu, off_w1, off_b1, partial, nslots, nfc + OC2, P, B, IC, H, W, OC1, pad1, m.PH1, m.PW1, st));
  cnn1_coef_kernel<<<P, 256, 0, st>>>(U, ldu, partial, nslots, clip, coef, s);
  cnn1_scale_kernel<<<dim3((ldu + 4095) / 4096, P), 256, 0, st>>>(U, ldu, coef, s);
  return finish_launch("cnn2_local_step", 11);
}

}  // extern "C"
