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
#include "tc_ptx.cuh"

namespace bsc {

static bool g_cnn_generic = false;
static bool g_cnn_wgrad_simt = false;

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
// (fpart != null: the logits arrive as nchunk partial sums per image, fpart[chunk][img][c], and are first summed
// in chunk order on top of the bias)
__global__ void cnn1_loss_kernel(float* __restrict__ dl, const int* __restrict__ y, float* __restrict__ loss,
                                 float* __restrict__ U, int ldu, int P, Cnn1Shape s, const float* __restrict__ fpart = nullptr,
                                 int nchunk = 0, const float* __restrict__ fbias = nullptr, float gsign = 1.f) {
  const int p = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (p >= P) return;
  float* sl = dl + (size_t)p * s.B * s.C;
  if (fpart) {
    const size_t PB = (size_t)P * s.B;
    for (int i = lane; i < s.B * s.C; i += 32) {
      const int b = i / s.C, c = i - b * s.C;
      float acc = 0.f;
      for (int k = 0; k < nchunk; k++) acc += fpart[((size_t)k * PB + (size_t)p * s.B + b) * s.C + c];
      sl[i] = acc + fbias[c];
    }
    __syncwarp();
  }
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
    U[(size_t)p * ldu + s.off_fb + lane] = gsign * g;   // unscaled gradient (or its negative); cnn1_scale finishes it
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
// negated = 1: the row already holds -grad, so a peer whose gradient is inside the clip radius (coef == 1, the common
// case) only gets its padding zeroed and the pass over its d floats is skipped
__global__ void __launch_bounds__(256) cnn1_scale_kernel(float* __restrict__ U, int ldu, const float* __restrict__ coef,
                                                         Cnn1Shape s, int negated = 0) {
  const int p = blockIdx.y;
  const float m = negated ? coef[p] : -coef[p];
  float* row = U + (size_t)p * ldu;
  const int k0 = blockIdx.x * 4096;
  if (negated && m == 1.f) {
    for (int k = max(k0, s.d) + threadIdx.x; k < min(ldu, k0 + 4096); k += blockDim.x) row[k] = 0.f;
    return;
  }
  for (int k = k0 + threadIdx.x; k < min(ldu, k0 + 4096); k += blockDim.x) row[k] = k < s.d ? row[k] * m : 0.f;
}

// ---------------------------------------------------------------------------------------------------
// Shape-specialised kernels for the two reference one-conv models (mnist_cnn: IC 1, K 5, pad 4, OC 16;
// cifar_cnn: IC 3, K 3, pad 3, OC 20; 10 classes).  Same math, same accumulation order and same scratch layout
// as the generic kernels above (which stay as the fallback for other shapes); what changes is where the data
// lives: tap and class loops are compile-time, so the accumulators are registers instead of local memory, the
// convolution stages its zero-padded image once per CTA and computes all output channels of a position from
// one register patch, the classifier keeps a slice of fc.weight in shared memory across many images, and the
// weight gradient reduces its 4 x taps accumulators with warp shuffles instead of one block reduction per tap.
// ---------------------------------------------------------------------------------------------------
constexpr int FCF_JCHUNK = 1024;   // fc.weight columns a CTA of cnn1_fc_fwd_t keeps in shared memory
constexpr int FCF_IMGS = 256;      // images per CTA of cnn1_fc_fwd_t
constexpr int WG_OCB = 4;          // output channels per CTA of cnn1_conv_wgrad_t

// grid PB: one image per CTA; thread -> output positions, all OC channels of a position from one register patch
template <int IC, int K, int PAD, int OC>
__global__ void __launch_bounds__(256) cnn1_conv_fwd_t(const float* __restrict__ w, const float* __restrict__ X,
                                                       float* __restrict__ act, Cnn1Shape s) {
  constexpr int TAPS = IC * K * K;
  extern __shared__ float sm[];
  float* sw = sm;                        // [OC][TAPS] then [OC] biases
  float* sx = sm + OC * TAPS + OC;       // [IC][HP][WP] zero-padded image
  const int HP = s.H + 2 * PAD, WP = s.W + 2 * PAD;
  const int img = blockIdx.x, tid = threadIdx.x;
  for (int i = tid; i < OC * TAPS + OC; i += 256) sw[i] = i < OC * TAPS ? w[s.off_cw + i] : w[s.off_cb + i - OC * TAPS];
  for (int i = tid; i < IC * HP * WP; i += 256) sx[i] = 0.f;
  __syncthreads();
  const float* x = X + (size_t)img * IC * s.H * s.W;
  for (int i = tid; i < IC * s.H * s.W; i += 256) {
    const int ix = i % s.W, iy = (i / s.W) % s.H, ic = i / (s.W * s.H);
    sx[(ic * HP + iy + PAD) * WP + ix + PAD] = x[i];
  }
  __syncthreads();
  const int npos = s.OH * s.OW;
  float* out = act + (size_t)img * OC * npos;
  for (int pos = tid; pos < npos; pos += 256) {
    const int oy = pos / s.OW, ox = pos - oy * s.OW;
    float patch[TAPS];
#pragma unroll
    for (int ic = 0; ic < IC; ic++)
#pragma unroll
      for (int ky = 0; ky < K; ky++)
#pragma unroll
        for (int kx = 0; kx < K; kx++) patch[(ic * K + ky) * K + kx] = sx[(ic * HP + oy + ky) * WP + ox + kx];
#pragma unroll 4
    for (int oc = 0; oc < OC; oc++) {
      float acc = sw[OC * TAPS + oc];
#pragma unroll
      for (int t = 0; t < TAPS; t++) acc = fmaf(sw[oc * TAPS + t], patch[t], acc);   // padded taps add exact zeros
      out[(size_t)oc * npos + pos] = acc > 0.f ? acc : 0.f;
    }
  }
}

// grid (ceil(J / FCF_JCHUNK), ceil(PB / FCF_IMGS)): partial logits of a J slice, one warp per image;
// part[chunk][img][c] (summed over the chunks in index order by cnn1_loss_kernel)
template <int C>
__global__ void __launch_bounds__(256) cnn1_fc_fwd_t(const float* __restrict__ w, const float* __restrict__ act,
                                                     float* __restrict__ part, Cnn1Shape s, int PB) {
  extern __shared__ float sW[];          // [C][FCF_JCHUNK]
  const int j0 = blockIdx.x * FCF_JCHUNK, jn = min(FCF_JCHUNK, s.J - j0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < C * FCF_JCHUNK; i += 256) {
    const int c = i / FCF_JCHUNK, jj = i - c * FCF_JCHUNK;
    sW[i] = jj < jn ? w[s.off_fw + (size_t)c * s.J + j0 + jj] : 0.f;
  }
  __syncthreads();
  const int i0 = blockIdx.y * FCF_IMGS, i1 = min(PB, i0 + FCF_IMGS);
  const int jn4 = jn >> 2;               // J and the chunk size are multiples of 4
  // two images per pass: every fc.weight value read from shared memory feeds two accumulators
  for (int img = i0 + warp; img < i1; img += 16) {
    const bool two = img + 8 < i1;
    const float4* a4 = reinterpret_cast<const float4*>(act + (size_t)img * s.J + j0);
    const float4* b4 = reinterpret_cast<const float4*>(act + (size_t)(two ? img + 8 : img) * s.J + j0);
    float acc[C], bcc[C];
#pragma unroll
    for (int c = 0; c < C; c++) acc[c] = bcc[c] = 0.f;
    for (int q = lane; q < jn4; q += 32) {
      const float4 av = a4[q], bv = b4[q];
#pragma unroll
      for (int c = 0; c < C; c++) {
        const float4 wv = *reinterpret_cast<const float4*>(sW + c * FCF_JCHUNK + 4 * q);
        acc[c] = fmaf(av.x, wv.x, acc[c]);
        acc[c] = fmaf(av.y, wv.y, acc[c]);
        acc[c] = fmaf(av.z, wv.z, acc[c]);
        acc[c] = fmaf(av.w, wv.w, acc[c]);
        bcc[c] = fmaf(bv.x, wv.x, bcc[c]);
        bcc[c] = fmaf(bv.y, wv.y, bcc[c]);
        bcc[c] = fmaf(bv.z, wv.z, bcc[c]);
        bcc[c] = fmaf(bv.w, wv.w, bcc[c]);
      }
    }
#pragma unroll
    for (int c = 0; c < C; c++) {
      acc[c] = warp_sum(acc[c]);
      bcc[c] = warp_sum(bcc[c]);
    }
    if (lane == 0) {
#pragma unroll
      for (int c = 0; c < C; c++) {
        part[((size_t)blockIdx.x * PB + img) * C + c] = acc[c];
        if (two) part[((size_t)blockIdx.x * PB + img + 8) * C + c] = bcc[c];
      }
    }
  }
}

// grid (ceil(J/256), P): as cnn1_fc_bwd_kernel with the class loops unrolled (accumulators in registers)
template <int C>
__global__ void __launch_bounds__(256) cnn1_fc_bwd_t(const float* __restrict__ w, float* __restrict__ act,
                                                     const float* __restrict__ dl, float* __restrict__ U, int ldu,
                                                     float* __restrict__ partial, int nslots, Cnn1Shape s) {
  __shared__ float sdl[32 * C];
  __shared__ float red[40];
  const int p = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
  for (int i = threadIdx.x; i < s.B * C; i += blockDim.x) sdl[i] = dl[(size_t)p * s.B * C + i];
  __syncthreads();
  float nrm = 0.f;
  if (j < s.J) {
    float wc[C], g[C];
#pragma unroll
    for (int c = 0; c < C; c++) {
      wc[c] = w[s.off_fw + (size_t)c * s.J + j];
      g[c] = 0.f;
    }
    float* a = act + (size_t)p * s.B * s.J + j;
    for (int b = 0; b < s.B; b++) {
      const float av = a[(size_t)b * s.J];
      float da = 0.f;
#pragma unroll
      for (int c = 0; c < C; c++) {
        const float d = sdl[b * C + c];
        g[c] = fmaf(d, av, g[c]);
        da = fmaf(d, wc[c], da);
      }
      a[(size_t)b * s.J] = av > 0.f ? da : 0.f;   // d(pre-activation)
    }
#pragma unroll
    for (int c = 0; c < C; c++) {
      U[(size_t)p * ldu + s.off_fw + (size_t)c * s.J + j] = -g[c];   // the update is minus the gradient
      nrm = fmaf(g[c], g[c], nrm);
    }
  }
  const float t = block_sum(nrm, red);
  if (threadIdx.x == 0) partial[(size_t)p * nslots + blockIdx.x] = t;
}

// grid (ceil(OC / WG_OCB), P): conv.weight / conv.bias gradients of WG_OCB output channels of one peer.  The zero-padded
// images of the whole minibatch are staged in shared memory once (B x IC x HP x WP floats); a thread walks the
// (image, position) pairs with the d(pre-activation) loads of WG_NB pairs in flight before the first is used, reads a
// position's patch once and feeds WG_OCB x TAPS register accumulators; per-channel partial norms go to slot0 + oc.
constexpr int WG_NB = 4;
template <int IC, int K, int PAD, int OC>
__global__ void __launch_bounds__(256) cnn1_conv_wgrad_t(const float* __restrict__ X, const float* __restrict__ dpre,
                                                         float* __restrict__ U, int ldu, float* __restrict__ partial,
                                                         int nslots, int slot0, Cnn1Shape s) {
  constexpr int TAPS = IC * K * K;
  extern __shared__ float sm[];
  const int HP = s.H + 2 * PAD, WP = s.W + 2 * PAD, IMG = IC * HP * WP;
  float* sx = sm;                                      // [B][IC][HP][WP]
  float* sred = sm + (size_t)s.B * IMG;                // [8 warps][WG_OCB][TAPS + 1]
  const int oc0 = blockIdx.x * WG_OCB, p = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int npos = s.OH * s.OW;
  float acc[WG_OCB][TAPS], accb[WG_OCB];
#pragma unroll
  for (int o = 0; o < WG_OCB; o++) {
    accb[o] = 0.f;
#pragma unroll
    for (int t = 0; t < TAPS; t++) acc[o][t] = 0.f;
  }
  for (int i = tid; i < s.B * IMG; i += 256) sx[i] = 0.f;
  __syncthreads();
  {
    const float* x = X + (size_t)p * s.B * IC * s.H * s.W;
    const int per = IC * s.H * s.W;
    for (int i = tid; i < s.B * per; i += 256) {
      const int b = i / per, r = i - b * per;
      const int ix = r % s.W, iy = (r / s.W) % s.H, ic = r / (s.W * s.H);
      sx[b * IMG + (ic * HP + iy + PAD) * WP + ix + PAD] = x[i];
    }
  }
  __syncthreads();
  const float* dp = dpre + (size_t)p * s.B * OC * npos;
  const int total = s.B * npos;
  for (int q0 = tid; q0 < total; q0 += 256 * WG_NB) {
    float d[WG_NB][WG_OCB];
    int xo[WG_NB];
#pragma unroll
    for (int u = 0; u < WG_NB; u++) {
      const int q = q0 + u * 256;
      const bool ok = q < total;
      const int b = ok ? q / npos : 0, pos = ok ? q - b * npos : 0;
      const int oy = pos / s.OW, ox = pos - oy * s.OW;
      xo[u] = b * IMG + oy * WP + ox;
#pragma unroll
      for (int o = 0; o < WG_OCB; o++)
        d[u][o] = (ok && oc0 + o < OC) ? __ldg(dp + ((size_t)b * OC + oc0 + o) * npos + pos) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < WG_NB; u++) {
      const float* xb = sx + xo[u];
#pragma unroll
      for (int ic = 0; ic < IC; ic++)
#pragma unroll
        for (int ky = 0; ky < K; ky++)
#pragma unroll
          for (int kx = 0; kx < K; kx++) {
            const float xv = xb[(ic * HP + ky) * WP + kx];
#pragma unroll
            for (int o = 0; o < WG_OCB; o++) acc[o][(ic * K + ky) * K + kx] = fmaf(d[u][o], xv, acc[o][(ic * K + ky) * K + kx]);
          }
#pragma unroll
      for (int o = 0; o < WG_OCB; o++) accb[o] += d[u][o];
    }
  }
  // fixed-shape reduction: lanes by shuffle, then the eight warps in index order
#pragma unroll
  for (int o = 0; o < WG_OCB; o++) {
#pragma unroll
    for (int t = 0; t < TAPS; t++) {
      const float v = warp_sum(acc[o][t]);
      if (lane == 0) sred[(warp * WG_OCB + o) * (TAPS + 1) + t] = v;
    }
    const float vb = warp_sum(accb[o]);
    if (lane == 0) sred[(warp * WG_OCB + o) * (TAPS + 1) + TAPS] = vb;
  }
  __syncthreads();
  if (tid < WG_OCB * (TAPS + 1)) {
    const int o = tid / (TAPS + 1), t = tid - o * (TAPS + 1);
    float g = 0.f;
    for (int wi = 0; wi < 8; wi++) g += sred[(wi * WG_OCB + o) * (TAPS + 1) + t];
    if (oc0 + o < OC) {
      if (t < TAPS) U[(size_t)p * ldu + s.off_cw + (oc0 + o) * TAPS + t] = -g;
      else U[(size_t)p * ldu + s.off_cb + oc0 + o] = -g;
    }
    sred[(0 * WG_OCB + o) * (TAPS + 1) + t] = g * g;    // warp 0's row is free now: squares for the channel norm
  }
  __syncthreads();
  if (tid < WG_OCB && oc0 + tid < OC) {
    float nrm = 0.f;
    for (int t = 0; t <= TAPS; t++) nrm += sred[(0 * WG_OCB + tid) * (TAPS + 1) + t];
    partial[(size_t)p * nslots + slot0 + oc0 + tid] = nrm;
  }
}

// thread -> one image: logits = bias + partial sums in chunk order, arg-max against the label (first maximum wins, as
// torch.argmax does), one atomicAdd per warp
template <int C>
__global__ void __launch_bounds__(256) cnn1_eval_count_t(const float* __restrict__ fpart, int nchunk, const float* __restrict__ fbias,
                                                         const int* __restrict__ y, int N, int* __restrict__ wrong) {
  const int img = blockIdx.x * blockDim.x + threadIdx.x;
  bool bad = false;
  if (img < N) {
    float best = -INFINITY;
    int arg = 0;
#pragma unroll
    for (int c = 0; c < C; c++) {
      float acc = 0.f;
      for (int k = 0; k < nchunk; k++) acc += fpart[((size_t)k * N + img) * C + c];
      acc += fbias[c];
      if (acc > best) {
        best = acc;
        arg = c;
      }
    }
    bad = arg != y[img];
  }
  const unsigned m = __ballot_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(wrong, __popc(m));
}

template <int IC, int K, int PAD, int OC, int C>
static int cnn1_eval_t(const float* w, const float* X, const int* y, int* wrong, float* act, float* fpart, int N,
                       const Cnn1Shape& s, cudaStream_t st) {
  constexpr int TAPS = IC * K * K;
  const int HP = s.H + 2 * PAD, WP = s.W + 2 * PAD;
  const int nchunk = (s.J + FCF_JCHUNK - 1) / FCF_JCHUNK;
  const size_t smem_fwd = (size_t)(OC * TAPS + OC + IC * HP * WP) * sizeof(float);
  const size_t smem_fc = (size_t)C * FCF_JCHUNK * sizeof(float);
  static bool configured = false;
  if (!configured) {
    BSC_CHECK(cudaFuncSetAttribute(cnn1_conv_fwd_t<IC, K, PAD, OC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    BSC_CHECK(cudaFuncSetAttribute(cnn1_fc_fwd_t<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    configured = true;
  }
  if (smem_fwd > 96 * 1024) return -1;
  BSC_CHECK(cudaMemsetAsync(wrong, 0, sizeof(int), st));
  cnn1_conv_fwd_t<IC, K, PAD, OC><<<N, 256, smem_fwd, st>>>(w, X, act, s);
  cnn1_fc_fwd_t<C><<<dim3(nchunk, (N + FCF_IMGS - 1) / FCF_IMGS), 256, smem_fc, st>>>(w, act, fpart, s, N);
  cnn1_eval_count_t<C><<<(N + 255) / 256, 256, 0, st>>>(fpart, nchunk, w + s.off_fb, y, N, wrong);
  return finish_launch("cnn1_eval", 3);
}

// ---------------------------------------------------------------------------------------------------
// The convolution's weight gradient on the tensor cores.  Per peer it is one small GEMM with a long reduction:
//   G[oc][t] = sum over the B*OH*OW (image, position) pairs of  D[oc][pos] * X[t][pos],
// D = d(pre-activation) (fp32, contiguous in pos: already K-major), X = im2col rows, one per tap t = (ic, ky, kx), plus
// one row of ones whose column is the bias gradient.  One CTA per peer: the zero-padded images of its minibatch are
// staged in shared memory once; per 64-position K block all threads write the bf16 hi / lo planes of both operands in
// the 128B-swizzled K-major layout (D straight from global with the next block's loads in flight, im2col out of the
// staged images), one thread issues the [128 x 32] += A * B^T UMMAs (hi*hi, hi*lo, lo*hi) into TMEM - the A tile is
// declared 128 rows tall, rows past OC alias the bytes behind it and produce lanes nobody reads - and warp 0 reads the
// OC result lanes at the end.  Two operand stages, recycled through tcgen05.commit -> mbarrier.
// ---------------------------------------------------------------------------------------------------
constexpr int WT_THREADS = 512;
constexpr int WT_NB = 32;                                  // padded N: taps + the bias column
constexpr int WT_AROWS = 24;                               // A rows staged per K block (OC <= 24): three swizzle atoms
constexpr int WT_A_BYTES = WT_AROWS * 128;
constexpr int WT_B_BYTES = WT_NB * 128;
constexpr int WT_STAGE = 2 * WT_A_BYTES + 2 * WT_B_BYTES;  // A_hi | A_lo | B_hi | B_lo
constexpr int WT_STAGES = 2;
constexpr int WT_TAIL = 6 * 1024;                          // what the last A_lo's 128-row window reaches past its stage

template <int IC, int K, int PAD, int OC>
__global__ void __launch_bounds__(WT_THREADS, 1) cnn1_conv_wgrad_tc(const float* __restrict__ X, const float* __restrict__ dpre,
                                                                    float* __restrict__ U, int ldu, float* __restrict__ partial,
                                                                    int nslots, int slot0, Cnn1Shape s) {
  constexpr int TAPS = IC * K * K;
  static_assert(TAPS + 1 <= WT_NB && OC <= WT_AROWS, "shape does not fit the tile");
  constexpr int A_ITEMS = OC * 16, B_ITEMS = (TAPS + 1) * 16, ITEMS = A_ITEMS + B_ITEMS;
  constexpr int PER = (ITEMS + WT_THREADS - 1) / WT_THREADS;
  extern __shared__ uint8_t wt_smem[];
  __shared__ __align__(8) unsigned long long sBar[WT_STAGES + 1];
  __shared__ uint32_t sTmem;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, p = blockIdx.x;
  const int HP = s.H + 2 * PAD, WP = s.W + 2 * PAD, IMG = IC * HP * WP;
  const int npos = s.OH * s.OW, total = s.B * npos, nkb = (total + 63) / 64;
  const uint32_t raw = smem_u32(wt_smem);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = wt_smem + (base - raw);
  float* sx = reinterpret_cast<float*>(gen + WT_STAGES * WT_STAGE + WT_TAIL);     // [B][IC][HP][WP]

  if (tid == 0) {
    for (int i = 0; i <= WT_STAGES; i++) mbar_init(smem_u32(&sBar[i]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sTmem)), "r"(32)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = tid; i < s.B * IMG; i += WT_THREADS) sx[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sTmem;
  {
    const float* x = X + (size_t)p * s.B * IC * s.H * s.W;
    const int per = IC * s.H * s.W;
    for (int i = tid; i < s.B * per; i += WT_THREADS) {
      const int b = i / per, r = i - b * per;
      const int ix = r % s.W, iy = (r / s.W) % s.H, ic = r / (s.W * s.H);
      sx[b * IMG + (ic * HP + iy + PAD) * WP + ix + PAD] = x[i];
    }
  }
  __syncthreads();

  // A thread owns the same PER (row, quad-of-positions) slots of every K block; only the positions move, by 64 per block.
  // kind 0: a row of D (row = output channel), kind 1: an im2col row (row = tap), kind 2: the row of ones, 3: nothing.
  const float* dp = dpre + (size_t)p * s.B * OC * npos;
  int kind[PER], soff[PER], pos[PER], img[PER], oy[PER], ox[PER];
  const float* arow[PER];
  int xoff[PER];
#pragma unroll
  for (int u = 0; u < PER; u++) {
    const int it = tid + u * WT_THREADS;
    const int q = it & 15;
    pos[u] = q << 2;                 // position inside the image, advanced below
    img[u] = 0;
    oy[u] = pos[u] / s.OW;
    ox[u] = pos[u] - oy[u] * s.OW;
    arow[u] = dp;
    xoff[u] = 0;
    if (it < A_ITEMS) {
      kind[u] = 0;
      soff[u] = (int)sw128_off(it >> 4, q, 0);
      arow[u] = dp + (size_t)(it >> 4) * npos;              // + img * OC * npos + pos
    } else if (it < ITEMS) {
      const int t = (it - A_ITEMS) >> 4;
      kind[u] = t == TAPS ? 2 : 1;
      soff[u] = 2 * WT_A_BYTES + (int)sw128_off(t, q, 0);
      const int tt = t < TAPS ? t : 0;
      const int ic = tt / (K * K), ky = (tt / K) % K, kx = tt % K;
      xoff[u] = (ic * HP + ky) * WP + kx;
    } else {
      kind[u] = 3;
      soff[u] = 0;
    }
  }
  const int d_oy = 64 / s.OW, d_ox = 64 - d_oy * s.OW;      // how 64 positions move (oy, ox)
  auto load_a = [&](float4* v) {                             // D values of the block the indices point at
#pragma unroll
    for (int u = 0; u < PER; u++) {
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (kind[u] == 0 && img[u] < s.B) v[u] = __ldg(reinterpret_cast<const float4*>(arow[u] + (size_t)img[u] * OC * npos + pos[u]));
    }
  };
  float4 cur[PER];
  load_a(cur);
  const uint32_t idesc = umma_idesc_bf16_128xN(WT_NB);
  for (int kb = 0; kb < nkb; kb++) {
    const int st = kb % WT_STAGES;
    const uint32_t so = (uint32_t)st * WT_STAGE;
    // im2col values of this block (shared memory), before the indices move on
    float4 xv[PER];
#pragma unroll
    for (int u = 0; u < PER; u++) {
      xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (img[u] < s.B) {
        if (kind[u] == 1) {
          const float* xr = sx + img[u] * IMG + xoff[u] + oy[u] * WP + ox[u];   // OW % 4 == 0: the four positions share a row
          xv[u] = make_float4(xr[0], xr[1], xr[2], xr[3]);
        } else if (kind[u] == 2) {
          xv[u] = make_float4(1.f, 1.f, 1.f, 1.f);             // the bias column: sum of D over the positions
        }
      }
    }
    // advance every slot by 64 positions, then start the next block's D loads (in flight during the conversion below)
#pragma unroll
    for (int u = 0; u < PER; u++) {
      pos[u] += 64;
      oy[u] += d_oy;
      ox[u] += d_ox;
      if (ox[u] >= s.OW) {
        ox[u] -= s.OW;
        oy[u]++;
      }
      if (pos[u] >= npos) {
        pos[u] -= npos;
        img[u]++;
        oy[u] = pos[u] / s.OW;
        ox[u] = pos[u] - oy[u] * s.OW;
      }
    }
    float4 nxt[PER];
    load_a(nxt);
    if (kb >= WT_STAGES) mbar_wait(smem_u32(&sBar[st]), (uint32_t)(kb / WT_STAGES - 1) & 1u);   // the MMAs that read this stage
#pragma unroll
    for (int u = 0; u < PER; u++) {
      if (kind[u] == 3) continue;
      uint2 hi, lo;
      split4(kind[u] == 0 ? cur[u] : xv[u], hi, lo);
      const int pl = kind[u] == 0 ? WT_A_BYTES : WT_B_BYTES;
      *reinterpret_cast<uint2*>(gen + so + soff[u]) = hi;
      *reinterpret_cast<uint2*>(gen + so + soff[u] + pl) = lo;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint64_t a_hi = umma_desc_sw128(base + so), a_lo = umma_desc_sw128(base + so + WT_A_BYTES);
      const uint64_t b_hi = umma_desc_sw128(base + so + 2 * WT_A_BYTES), b_lo = umma_desc_sw128(base + so + 2 * WT_A_BYTES + WT_B_BYTES);
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint64_t off = (uint64_t)((k * 16 * 2) >> 4);
        umma_bf16(tmem, a_hi + off, b_hi + off, idesc, (kb > 0 || k > 0) ? 1u : 0u);
        umma_bf16(tmem, a_hi + off, b_lo + off, idesc, 1u);
        umma_bf16(tmem, a_lo + off, b_hi + off, idesc, 1u);
      }
      umma_commit(smem_u32(&sBar[st]));
      if (kb == nkb - 1) umma_commit(smem_u32(&sBar[WT_STAGES]));
    }
#pragma unroll
    for (int u = 0; u < PER; u++) cur[u] = nxt[u];
  }
  // ---- epilogue: lane = output channel, column = tap (column TAPS = bias) ------------------------------------
  if (warp == 0) {
    mbar_wait(smem_u32(&sBar[WT_STAGES]), 0);
    tc_fence_after();
    uint32_t v[32];
    tmem_ld32(tmem, v);
    if (lane < OC) {
      float nrm = 0.f;
      float* out = U + (size_t)p * ldu;
#pragma unroll
      for (int t = 0; t < TAPS; t++) {
        const float g = __uint_as_float(v[t]);
        out[s.off_cw + lane * TAPS + t] = -g;
        nrm = fmaf(g, g, nrm);
      }
      const float gb = __uint_as_float(v[TAPS]);
      out[s.off_cb + lane] = -gb;
      partial[(size_t)p * nslots + slot0 + lane] = fmaf(gb, gb, nrm);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(32) : "memory");
  }
}

template <int IC, int K, int PAD, int OC, int C>
static int cnn1_local_step_t(const float* w, const float* X, const int* y, float* U, float* loss, float* act, float* dl,
                             float* partial, float* coef, float* fpart, int P, int ldu, float clip, const Cnn1Shape& s,
                             cudaStream_t st) {
  constexpr int TAPS = IC * K * K;
  const int PB = P * s.B, HP = s.H + 2 * PAD, WP = s.W + 2 * PAD;
  const int nfc = (s.J + 255) / 256, nslots = nfc + OC;
  const int nchunk = (s.J + FCF_JCHUNK - 1) / FCF_JCHUNK;
  const size_t smem_fwd = (size_t)(OC * TAPS + OC + IC * HP * WP) * sizeof(float);
  const size_t smem_fc = (size_t)C * FCF_JCHUNK * sizeof(float);
  const size_t smem_wg = ((size_t)s.B * IC * HP * WP + 8 * WG_OCB * (TAPS + 1)) * sizeof(float);
  static bool configured = false;
  if (!configured) {
    BSC_CHECK(cudaFuncSetAttribute(cnn1_conv_fwd_t<IC, K, PAD, OC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    BSC_CHECK(cudaFuncSetAttribute(cnn1_fc_fwd_t<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    BSC_CHECK(cudaFuncSetAttribute(cnn1_conv_wgrad_t<IC, K, PAD, OC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    configured = true;
  }
  if (smem_fwd > 96 * 1024 || smem_wg > 220 * 1024) return -1;
  cnn1_conv_fwd_t<IC, K, PAD, OC><<<PB, 256, smem_fwd, st>>>(w, X, act, s);
  cnn1_fc_fwd_t<C><<<dim3(nchunk, (PB + FCF_IMGS - 1) / FCF_IMGS), 256, smem_fc, st>>>(w, act, fpart, s, PB);
  cnn1_loss_kernel<<<(P + 3) / 4, 128, 0, st>>>(dl, y, loss, U, ldu, P, s, fpart, nchunk, w + s.off_fb, -1.f);
  cnn1_fc_bwd_t<C><<<dim3(nfc, P), 256, 0, st>>>(w, act, dl, U, ldu, partial, nslots, s);
  const size_t smem_tc = 1024 + (size_t)WT_STAGES * WT_STAGE + WT_TAIL + (size_t)s.B * IC * HP * WP * sizeof(float);
  if (!g_cnn_wgrad_simt && smem_tc <= 224 * 1024 && (s.OW & 3) == 0) {   // 227 KB per CTA minus the kernel's static shared memory
    static bool configured_tc = false;
    if (!configured_tc) {
      BSC_CHECK(cudaFuncSetAttribute(cnn1_conv_wgrad_tc<IC, K, PAD, OC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
      configured_tc = true;
    }
    cnn1_conv_wgrad_tc<IC, K, PAD, OC><<<P, WT_THREADS, smem_tc, st>>>(X, act, U, ldu, partial, nslots, nfc, s);
  } else {
    cnn1_conv_wgrad_t<IC, K, PAD, OC><<<dim3((OC + WG_OCB - 1) / WG_OCB, P), 256, smem_wg, st>>>(X, act, U, ldu, partial, nslots,
                                                                                               nfc, s);
  }
  cnn1_coef_kernel<<<P, 256, 0, st>>>(U, ldu, partial, nslots, clip, coef, s);
  cnn1_scale_kernel<<<dim3((ldu + 4095) / 4096, P), 256, 0, st>>>(U, ldu, coef, s, 1);
  return finish_launch("cnn1_local_step", 7);
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

// test hook: 1 = keep the one-conv models on the generic kernels (the shape-specialised ones are compared against them)
void bsc_cnn_force_generic(int on) { g_cnn_generic = on != 0; }
// test hook: 1 = conv weight gradient on the SIMT kernel instead of the tensor-core one
void bsc_cnn_force_wgrad_simt(int on) { g_cnn_wgrad_simt = on != 0; }

// scratch: act [P*B*J] floats, dl [P*B*C], partial [P*nslots] with nslots = ceil(J/256) + OC, coef [P],
// fpart [ceil(J/1024)][P*B][C]
long long bsc_cnn1_scratch_floats(int P, int B, int IC, int OC, int K, int pad, int H, int W, int C) {
  const int OH = H + 2 * pad - K + 1, OW = W + 2 * pad - K + 1;
  const long long J = (long long)OC * OH * OW;
  const long long nslots = (J + 255) / 256 + OC;
  const long long nchunk = (J + FCF_JCHUNK - 1) / FCF_JCHUNK;   // partial logits of the shape-specialised classifier
  return (long long)P * B * J + (long long)P * B * C + (long long)P * nslots + P + nchunk * P * B * C;
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
  float* fpart = coef + P;
  if (C == 10 && (s.J & 3) == 0 && !g_cnn_generic) {
    if (IC == 3 && K == 3 && pad == 3 && OC == 20)
      return cnn1_local_step_t<3, 3, 3, 20, 10>(w, X, y, U, loss, act, dl, partial, coef, fpart, P, ldu, clip, s, st);
    if (IC == 1 && K == 5 && pad == 4 && OC == 16)
      return cnn1_local_step_t<1, 5, 4, 16, 10>(w, X, y, U, loss, act, dl, partial, coef, fpart, P, ldu, clip, s, st);
  }
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

// Error count of one weight vector over N images with the one-conv model (K14 for the CNN families; the reference
// runs the torch module, ML/Pytorch/client.py:136-172).  wrong: int32[1] out (zeroed here); scratch:
// bsc_cnn1_eval_scratch_floats floats.  Only the two reference shapes with 10 classes are supported (-1 otherwise: the
// caller falls back to the framework's convolution).
long long bsc_cnn1_eval_scratch_floats(int N, int IC, int OC, int K, int pad, int H, int W, int C) {
  const int OH = H + 2 * pad - K + 1, OW = W + 2 * pad - K + 1;
  const long long J = (long long)OC * OH * OW;
  return (long long)N * J + ((J + FCF_JCHUNK - 1) / FCF_JCHUNK) * N * C;
}

int bsc_cnn1_eval_errors(const float* w, const float* X, const int* y, int* wrong, float* scratch, int N, int IC, int OC,
                         int K, int pad, int H, int W, int C, void* stream) {
  Cnn1Shape s;
  s.IC = IC; s.OC = OC; s.K = K; s.pad = pad; s.H = H; s.W = W;
  s.OH = H + 2 * pad - K + 1;
  s.OW = W + 2 * pad - K + 1;
  s.J = OC * s.OH * s.OW;
  s.C = C;
  s.B = 1;
  s.off_cw = 0;
  s.off_cb = OC * IC * K * K;
  s.off_fw = s.off_cb + OC;
  s.off_fb = s.off_fw + C * s.J;
  s.d = s.off_fb + C;
  float* act = scratch;
  float* fpart = act + (size_t)N * s.J;
  cudaStream_t st = (cudaStream_t)stream;
  if (C == 10 && (s.J & 3) == 0) {
    if (IC == 3 && K == 3 && pad == 3 && OC == 20) return cnn1_eval_t<3, 3, 3, 20, 10>(w, X, y, wrong, act, fpart, N, s, st);
    if (IC == 1 && K == 5 && pad == 4 && OC == 16) return cnn1_eval_t<1, 5, 4, 16, 10>(w, X, y, wrong, act, fpart, N, s, st);
  }
  return -1;
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
