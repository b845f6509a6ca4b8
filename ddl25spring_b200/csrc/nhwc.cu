// NHWC companions of the tap-GEMM convolutions (conv_tcgen05.cu): everything between two convolutions of ResNet-18 /
// DCGAN stays channels-last, so each op below is one coalesced pass over [rows = N*H*W, C] fp32.
//
//   bn_stats      per-channel sum / sum of squares in ordered partials; the last CTA of a channel group turns them into
//                 mean / rstd and updates the running statistics (one launch, deterministic)
//   bn_apply      y = act((x - mean) * rstd * gamma + beta [+ residual])         (BatchNorm + ReLU/LeakyReLU + shortcut add fused)
//   bn_bwd_reduce dz = dy * [y > 0];  dbeta = sum dz, dgamma = sum dz * xhat      (same partial scheme)
//   bn_bwd_apply  dx = gamma * rstd * (dz - dbeta/R - xhat * dgamma/R);  dres = dz
//   phase_split   the four even/odd sub-images of x (input of a stride-2 convolution)
//   maxpool3x3s2  forward (+ argmax code) / backward (gather form, no atomics)
//   avgpool       global average pool forward / backward
//   im2col_stem   7x7 stride-2 pad-3 patches of the NCHW network input (Cin = 3 padded to 4): the stem is one plain GEMM
//
// Replaces (reference): ATen batch_norm / relu / add / max_pool2d / adaptive_avg_pool2d of the north-star ResNet-18
// (SURVEY §2.7, §8 K11/K12).
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

constexpr int kBnThreads = 256;

// Column reductions over [R, C]: a CTA owns 4*L channels (L = min(C/4, 32) float4 lanes per row) and a slice of the rows;
// thread (lx, ry) accumulates rows ry, ry + RL, ... of channel quad lx with four independent chains (loads in flight).
struct ColGeom {
  int L, RL, lx, ry, c;       // c = first channel of this thread's quad
  int64_t r0, r1;
};
__device__ __forceinline__ ColGeom col_geom(int64_t R, int C, int n_split) {
  ColGeom g;
  g.L = min(C >> 2, 32);
  g.RL = kBnThreads / g.L;
  g.lx = threadIdx.x % g.L;
  g.ry = threadIdx.x / g.L;
  g.c = (blockIdx.x * g.L + g.lx) * 4;
  const int64_t chunk = (R + n_split - 1) / n_split;
  g.r0 = blockIdx.y * chunk;
  g.r1 = min(R, g.r0 + chunk);
  return g;
}

__device__ __forceinline__ void add4(float4& a, const float4& v) { a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; }

// partial[(split * 2 + which) * C + c]; f(row, c, a, b) accumulates one row's contribution into the float4 pair (a, b)
template <typename F>
__device__ __forceinline__ void column_partials(int64_t R, int C, int n_split, float* __restrict__ partial, F f) {
  __shared__ float4 sa[kBnThreads], sb[kBnThreads];
  const ColGeom g = col_geom(R, C, n_split);
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
  if (g.c < C) {
    float4 a1 = a, b1 = a, a2 = a, b2 = a, a3 = a, b3 = a;
    int64_t r = g.r0 + g.ry;
    const int st = g.RL;
#pragma unroll 2
    for (; r + 3 * st < g.r1; r += 4 * st) { f(r, g.c, a, b); f(r + st, g.c, a1, b1); f(r + 2 * st, g.c, a2, b2); f(r + 3 * st, g.c, a3, b3); }
    for (; r < g.r1; r += st) f(r, g.c, a, b);
    add4(a2, a3); add4(a, a1); add4(a, a2);
    add4(b2, b3); add4(b, b1); add4(b, b2);
  }
  sa[threadIdx.x] = a; sb[threadIdx.x] = b;
  __syncthreads();
  if (g.ry == 0 && g.c < C) {
    for (int k = 1; k < g.RL; ++k) { add4(a, sa[k * g.L + g.lx]); add4(b, sb[k * g.L + g.lx]); }
    *reinterpret_cast<float4*>(partial + ((int64_t)blockIdx.y * 2 + 0) * C + g.c) = a;
    *reinterpret_cast<float4*>(partial + ((int64_t)blockIdx.y * 2 + 1) * C + g.c) = b;
  }
}

// true in exactly one CTA of the channel group: the one that finished last
__device__ __forceinline__ bool last_block_of_group(unsigned* counters) {
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned done = atomicAdd(&counters[blockIdx.x], 1u);
    is_last = (done == gridDim.y - 1);
    if (is_last) counters[blockIdx.x] = 0;      // ready for the next launch
  }
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}

// Sum the per-split partials of this CTA's channels in a fixed order: row lane ry adds splits ry, ry + RL, ...; lane 0 adds
// the RL lane sums.  Valid in the threads with ry == 0: a[i], b[i] for channels c .. c+3.
__device__ __forceinline__ bool combine_partials(const float* __restrict__ partial, int64_t R, int C, int n_split, double (&a)[4],
                                                 double (&b)[4], int& c_out) {
  __shared__ double da[kBnThreads][4], db[kBnThreads][4];
  const ColGeom g = col_geom(R, C, n_split);
#pragma unroll
  for (int i = 0; i < 4; ++i) { a[i] = 0.0; b[i] = 0.0; }
  if (g.c < C)
    for (int k = g.ry; k < n_split; k += g.RL) {
      const float4 va = __ldcg(reinterpret_cast<const float4*>(partial + ((int64_t)k * 2 + 0) * C + g.c));
      const float4 vb = __ldcg(reinterpret_cast<const float4*>(partial + ((int64_t)k * 2 + 1) * C + g.c));
      a[0] += va.x; a[1] += va.y; a[2] += va.z; a[3] += va.w;
      b[0] += vb.x; b[1] += vb.y; b[2] += vb.z; b[3] += vb.w;
    }
#pragma unroll
  for (int i = 0; i < 4; ++i) { da[threadIdx.x][i] = a[i]; db[threadIdx.x][i] = b[i]; }
  __syncthreads();
  c_out = g.c;
  if (g.ry != 0 || g.c >= C) return false;
  for (int k = 1; k < g.RL; ++k)
#pragma unroll
    for (int i = 0; i < 4; ++i) { a[i] += da[k * g.L + g.lx][i]; b[i] += db[k * g.L + g.lx][i]; }
  return true;
}

__global__ void __launch_bounds__(kBnThreads) bn_stats_kernel(const float* __restrict__ x, int64_t R, int C, float* __restrict__ partial,
                                                              unsigned* __restrict__ counters, float* __restrict__ mean,
                                                              float* __restrict__ rstd, float* __restrict__ run_mean,
                                                              float* __restrict__ run_var, long long* __restrict__ n_tracked, float eps,
                                                              float momentum) {
  pdl_launch_dependents();
  pdl_wait();
  column_partials(R, C, gridDim.y, partial, [&](int64_t r, int c, float4& a, float4& b) {
    const float4 v = *reinterpret_cast<const float4*>(x + r * C + c);
    add4(a, v);
    b.x = fmaf(v.x, v.x, b.x); b.y = fmaf(v.y, v.y, b.y); b.z = fmaf(v.z, v.z, b.z); b.w = fmaf(v.w, v.w, b.w);
  });
  if (!last_block_of_group(counters)) return;
  double s[4], q[4];
  int c;
  if (!combine_partials(partial, R, C, gridDim.y, s, q, c)) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const double m = s[i] / (double)R;
    double var = q[i] / (double)R - m * m;
    if (var < 0.0) var = 0.0;
    mean[c + i] = (float)m;
    rstd[c + i] = (float)(1.0 / sqrt(var + (double)eps));
    if (run_mean) {
      const double unbiased = R > 1 ? var * (double)R / (double)(R - 1) : var;
      run_mean[c + i] = (1.f - momentum) * run_mean[c + i] + momentum * (float)m;
      run_var[c + i] = (1.f - momentum) * run_var[c + i] + momentum * (float)unbiased;
    }
  }
  if (n_tracked && c == 0) *n_tracked += 1;
}

__global__ void __launch_bounds__(256) bn_apply_kernel(const float4* __restrict__ x, const float4* __restrict__ res, float4* __restrict__ y,
                                                       const float* __restrict__ mean, const float* __restrict__ rstd,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta, int64_t total4,
                                                       int C, int relu, float slope) {
  pdl_launch_dependents();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)((i * 4) % C);
    const float4 v = x[i];
    const float4 m = *reinterpret_cast<const float4*>(mean + c), s = *reinterpret_cast<const float4*>(rstd + c);
    const float4 g = *reinterpret_cast<const float4*>(gamma + c), b = *reinterpret_cast<const float4*>(beta + c);
    float4 o;
    o.x = (v.x - m.x) * s.x * g.x + b.x; o.y = (v.y - m.y) * s.y * g.y + b.y;
    o.z = (v.z - m.z) * s.z * g.z + b.z; o.w = (v.w - m.w) * s.w * g.w + b.w;
    if (res) { const float4 r = res[i]; o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w; }
    if (relu == 1) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
    else if (relu == 2) {
      o.x = o.x > 0.f ? o.x : o.x * slope; o.y = o.y > 0.f ? o.y : o.y * slope;
      o.z = o.z > 0.f ? o.z : o.z * slope; o.w = o.w > 0.f ? o.w : o.w * slope;
    }
    y[i] = o;
  }
}

__global__ void __launch_bounds__(kBnThreads) bn_bwd_reduce_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                                   const float* __restrict__ dy, int64_t R, int C,
                                                                   const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                   float* __restrict__ partial, unsigned* __restrict__ counters,
                                                                   float* __restrict__ dgamma, float* __restrict__ dbeta, int relu,
                                                                   float slope) {
  pdl_launch_dependents();
  pdl_wait();
  const ColGeom g0 = col_geom(R, C, gridDim.y);
  float4 m = make_float4(0.f, 0.f, 0.f, 0.f), s = m;
  if (g0.c < C) { m = *reinterpret_cast<const float4*>(mean + g0.c); s = *reinterpret_cast<const float4*>(rstd + g0.c); }
  column_partials(R, C, gridDim.y, partial, [&](int64_t r, int c, float4& a, float4& b) {
    float4 g = *reinterpret_cast<const float4*>(dy + r * C + c);
    if (relu) {
      const float neg = relu == 2 ? slope : 0.f;      // leaky: y and the pre-activation have the same sign
      const float4 o = *reinterpret_cast<const float4*>(y + r * C + c);
      if (!(o.x > 0.f)) g.x *= neg;
      if (!(o.y > 0.f)) g.y *= neg;
      if (!(o.z > 0.f)) g.z *= neg;
      if (!(o.w > 0.f)) g.w *= neg;
    }
    const float4 v = *reinterpret_cast<const float4*>(x + r * C + c);
    add4(a, g);
    b.x = fmaf(g.x, (v.x - m.x) * s.x, b.x); b.y = fmaf(g.y, (v.y - m.y) * s.y, b.y);
    b.z = fmaf(g.z, (v.z - m.z) * s.z, b.z); b.w = fmaf(g.w, (v.w - m.w) * s.w, b.w);
  });
  if (!last_block_of_group(counters)) return;
  double a[4], b[4];
  int c;
  if (!combine_partials(partial, R, C, gridDim.y, a, b, c)) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) { dbeta[c + i] = (float)a[i]; dgamma[c + i] = (float)b[i]; }
}

__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const float4* __restrict__ x, const float4* __restrict__ y,
                                                           const float4* __restrict__ dy, float4* __restrict__ dx, float4* __restrict__ dres,
                                                           const float* __restrict__ mean, const float* __restrict__ rstd,
                                                           const float* __restrict__ gamma, const float* __restrict__ dgamma,
                                                           const float* __restrict__ dbeta, int64_t total4, int C, float inv_rows, int relu,
                                                           float slope) {
  pdl_launch_dependents();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)((i * 4) % C);
    float4 g = dy[i];
    if (relu) {
      const float neg = relu == 2 ? slope : 0.f;
      const float4 o = y[i];
      if (!(o.x > 0.f)) g.x *= neg;
      if (!(o.y > 0.f)) g.y *= neg;
      if (!(o.z > 0.f)) g.z *= neg;
      if (!(o.w > 0.f)) g.w *= neg;
    }
    if (dres) dres[i] = g;
    const float4 v = x[i];
    const float4 m = *reinterpret_cast<const float4*>(mean + c), s = *reinterpret_cast<const float4*>(rstd + c);
    const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
    const float4 dg = *reinterpret_cast<const float4*>(dgamma + c), db = *reinterpret_cast<const float4*>(dbeta + c);
    float4 o;
    o.x = ga.x * s.x * (g.x - db.x * inv_rows - (v.x - m.x) * s.x * dg.x * inv_rows);
    o.y = ga.y * s.y * (g.y - db.y * inv_rows - (v.y - m.y) * s.y * dg.y * inv_rows);
    o.z = ga.z * s.z * (g.z - db.z * inv_rows - (v.z - m.z) * s.z * dg.z * inv_rows);
    o.w = ga.w * s.w * (g.w - db.w * inv_rows - (v.w - m.w) * s.w * dg.w * inv_rows);
    dx[i] = o;
  }
}

// x [N, H, W, C] -> ph [4][N, H/2, W/2, C], ph[2a+b][n,i,j,:] = x[n, 2i+a, 2j+b, :]
__global__ void __launch_bounds__(256) phase_split_kernel(const float4* __restrict__ x, float4* __restrict__ ph, int N, int H, int W, int C4) {
  pdl_launch_dependents();
  pdl_wait();
  const int Ho = H / 2, Wo = W / 2;
  const int64_t per = (int64_t)N * Ho * Wo * C4, total = 4 * per;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int phase = (int)(e / per);
    int64_t r = e % per;
    const int c = (int)(r % C4); r /= C4;
    const int j = (int)(r % Wo); r /= Wo;
    const int i = (int)(r % Ho);
    const int n = (int)(r / Ho);
    ph[e] = x[(((int64_t)n * H + 2 * i + (phase >> 1)) * W + 2 * j + (phase & 1)) * C4 + c];
  }
}

// 3x3 stride-2 pad-1 max pool, NHWC, four channels per thread.  code = r*3+s of the first maximum (row-major scan, like torch).
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const float4* __restrict__ x, float4* __restrict__ y, uchar4* __restrict__ code,
                                                          int N, int H, int W, int C4, int Ho, int Wo) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t total = (int64_t)N * Ho * Wo * C4;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    const int c = (int)(r % C4); r /= C4;
    const int j = (int)(r % Wo); r /= Wo;
    const int i = (int)(r % Ho);
    const int n = (int)(r / Ho);
    float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    uchar4 bc = make_uchar4(4, 4, 4, 4);
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        const int h = 2 * i + a - 1, w = 2 * j + b - 1;
        if (h < 0 || h >= H || w < 0 || w >= W) continue;
        const float4 v = x[(((int64_t)n * H + h) * W + w) * C4 + c];
        const unsigned char t = (unsigned char)(a * 3 + b);
        if (v.x > best.x) { best.x = v.x; bc.x = t; }
        if (v.y > best.y) { best.y = v.y; bc.y = t; }
        if (v.z > best.z) { best.z = v.z; bc.z = t; }
        if (v.w > best.w) { best.w = v.w; bc.w = t; }
      }
    y[e] = best;
    code[e] = bc;
  }
}

__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const float4* __restrict__ dy, const uchar4* __restrict__ code,
                                                          float4* __restrict__ dx, int N, int H, int W, int C4, int Ho, int Wo) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t total = (int64_t)N * H * W * C4;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    const int c = (int)(r % C4); r /= C4;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int n = (int)(r / H);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    // windows i with 2i-1 <= h <= 2i+1
    for (int i = h / 2; i <= (h + 1) / 2; ++i) {
      if (i >= Ho) continue;
      const int a = h - (2 * i - 1);
      for (int j = w / 2; j <= (w + 1) / 2; ++j) {
        if (j >= Wo) continue;
        const unsigned char t = (unsigned char)(a * 3 + (w - (2 * j - 1)));
        const int64_t o = (((int64_t)n * Ho + i) * Wo + j) * C4 + c;
        const uchar4 cd = code[o];
        const float4 g = dy[o];
        if (cd.x == t) acc.x += g.x;
        if (cd.y == t) acc.y += g.y;
        if (cd.z == t) acc.z += g.z;
        if (cd.w == t) acc.w += g.w;
      }
    }
    dx[e] = acc;
  }
}

// y[n, c] = mean_s x[n, s, c]
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int S, int C) {
  pdl_launch_dependents();
  pdl_wait();
  const int n = blockIdx.x;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f;
    for (int s = 0; s < S; ++s) a += x[((int64_t)n * S + s) * C + c];
    y[(int64_t)n * C + c] = a / (float)S;
  }
}
__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int64_t total, int S, int C) {
  pdl_launch_dependents();
  pdl_wait();
  const float inv = 1.f / (float)S;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(e % C);
    const int64_t n = e / ((int64_t)S * C);
    dx[e] = dy[n * C + c] * inv;
  }
}

// col[(n,i,j), (r*7+s)*4 + ci] = x[n, ci, 2i+r-3, 2j+s-3] (NCHW input, ci < Cin <= 4; zero elsewhere); row pitch 224 floats
__global__ void __launch_bounds__(256) im2col_stem_kernel(const float* __restrict__ x, float4* __restrict__ col, int N, int Cin, int H, int W,
                                                          int Ho, int Wo) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t total = (int64_t)N * Ho * Wo * 56;     // 49 taps + 7 zero float4s: rows of 224 floats (a multiple of 32)
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    const int t = (int)(r % 56); r /= 56;
    const int j = (int)(r % Wo); r /= Wo;
    const int i = (int)(r % Ho);
    const int n = (int)(r / Ho);
    const int h = 2 * i + t / 7 - 3, w = 2 * j + t % 7 - 3;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (t < 49 && h >= 0 && h < H && w >= 0 && w < W)
      for (int ci = 0; ci < Cin; ++ci) v[ci] = x[(((int64_t)n * Cin + ci) * H + h) * W + w];
    col[e] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

int grid_for(int64_t work, int threads) {
  int64_t b = (work + threads - 1) / threads;
  if (b > kNumSMs * 16) b = kNumSMs * 16;
  return (int)(b < 1 ? 1 : b);
}

int bn_groups(int C) { return (C / 4 + 31) / 32; }     // CTAs along the channel axis (128 channels each)

int bn_splits(int64_t R, int C) {
  const int groups = bn_groups(C);
  int64_t s = (2 * kNumSMs + groups - 1) / groups;       // about two CTAs per SM ...
  if (s > (R + 63) / 64) s = (R + 63) / 64;              // ... of at least 64 rows each (latency-bound: short chains, many CTAs)
  if (s > 128) s = 128;
  return (int)(s < 1 ? 1 : s);
}

}  // namespace

extern "C" {

// workspace: partial needs ddl_bn_nhwc_ws_floats(R, C) floats; counters: (C+31)/32 zero-initialised unsigned (left zero).
int64_t ddl_bn_nhwc_ws_floats(int64_t R, int C) { return (int64_t)bn_splits(R, C) * 2 * C; }

int ddl_bn_nhwc_stats(const float* x, int64_t R, int C, float* partial, unsigned* counters, float* mean, float* rstd, float* run_mean,
                      float* run_var, long long* n_tracked, float eps, float momentum, void* stream) {
  DDL_REQUIRE(R > 0 && C > 0 && C % 4 == 0 && (C < 128 || C % 128 == 0) && 256 % (C < 128 ? C / 4 : 32) == 0,
              "bn_nhwc: need rows > 0 and C a multiple of 4 that divides 128 or is a multiple of 128 (C=%d)", C);
  dim3 grid(bn_groups(C), bn_splits(R, C));
  DDL_CUDA_CHECK(ddl_launch_pdl(bn_stats_kernel, dim3(grid), dim3(kBnThreads), 0, (cudaStream_t)stream, x, R, C, partial, counters, mean, rstd, run_mean, run_var, n_tracked, eps, momentum));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_bn_nhwc_apply(const float* x, const float* res, float* y, const float* mean, const float* rstd, const float* gamma,
                      const float* beta, int64_t R, int C, int relu, float slope, void* stream) {
  DDL_REQUIRE(C % 4 == 0, "bn_nhwc: C %% 4 != 0");
  const int64_t total4 = R * C / 4;
  DDL_CUDA_CHECK(ddl_launch_pdl(bn_apply_kernel, dim3(grid_for(total4, 256)), dim3(256), 0, (cudaStream_t)stream, 
      reinterpret_cast<const float4*>(x), reinterpret_cast<const float4*>(res), reinterpret_cast<float4*>(y), mean, rstd, gamma, beta,
      total4, C, relu, slope));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_bn_nhwc_bwd(const float* x, const float* y, const float* dy, float* dx, float* dres, const float* mean, const float* rstd,
                    const float* gamma, float* dgamma, float* dbeta, float* partial, unsigned* counters, int64_t R, int C, int relu,
                    float slope, void* stream) {
  DDL_REQUIRE(R > 0 && C > 0 && C % 4 == 0 && (C < 128 || C % 128 == 0) && 256 % (C < 128 ? C / 4 : 32) == 0,
              "bn_nhwc: need rows > 0 and C a multiple of 4 that divides 128 or is a multiple of 128 (C=%d)", C);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(bn_groups(C), bn_splits(R, C));
  DDL_CUDA_CHECK(ddl_launch_pdl(bn_bwd_reduce_kernel, dim3(grid), dim3(kBnThreads), 0, st, x, y, dy, R, C, mean, rstd, partial, counters, dgamma, dbeta, relu, slope));
  DDL_LAUNCH_CHECK();
  const int64_t total4 = R * C / 4;
  DDL_CUDA_CHECK(ddl_launch_pdl(bn_bwd_apply_kernel, dim3(grid_for(total4, 256)), dim3(256), 0, st, 
      reinterpret_cast<const float4*>(x), reinterpret_cast<const float4*>(y), reinterpret_cast<const float4*>(dy),
      reinterpret_cast<float4*>(dx), reinterpret_cast<float4*>(dres), mean, rstd, gamma, dgamma, dbeta, total4, C, 1.f / (float)R, relu, slope));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_phase_split(const float* x, float* ph, int N, int H, int W, int C, void* stream) {
  DDL_REQUIRE(H % 2 == 0 && W % 2 == 0 && C % 4 == 0, "phase_split: H, W must be even and C %% 4 == 0");
  const int64_t total = (int64_t)N * H * W * (C / 4);
  DDL_CUDA_CHECK(ddl_launch_pdl(phase_split_kernel, dim3(grid_for(total, 256)), dim3(256), 0, (cudaStream_t)stream, reinterpret_cast<const float4*>(x),
                                                                            reinterpret_cast<float4*>(ph), N, H, W, C / 4));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_maxpool3x3s2_fwd(const float* x, float* y, unsigned char* code, int N, int H, int W, int C, void* stream) {
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
  DDL_REQUIRE(C % 4 == 0, "maxpool: C %% 4 != 0");
  DDL_CUDA_CHECK(ddl_launch_pdl(maxpool_fwd_kernel, dim3(grid_for((int64_t)N * Ho * Wo * (C / 4), 256)), dim3(256), 0, (cudaStream_t)stream, 
      reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(y), reinterpret_cast<uchar4*>(code), N, H, W, C / 4, Ho, Wo));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_maxpool3x3s2_bwd(const float* dy, const unsigned char* code, float* dx, int N, int H, int W, int C, void* stream) {
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
  DDL_REQUIRE(C % 4 == 0, "maxpool: C %% 4 != 0");
  DDL_CUDA_CHECK(ddl_launch_pdl(maxpool_bwd_kernel, dim3(grid_for((int64_t)N * H * W * (C / 4), 256)), dim3(256), 0, (cudaStream_t)stream, 
      reinterpret_cast<const float4*>(dy), reinterpret_cast<const uchar4*>(code), reinterpret_cast<float4*>(dx), N, H, W, C / 4, Ho, Wo));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_avgpool_nhwc_fwd(const float* x, float* y, int N, int S, int C, void* stream) {
  DDL_CUDA_CHECK(ddl_launch_pdl(avgpool_fwd_kernel, dim3(N), dim3(256), 0, (cudaStream_t)stream, x, y, S, C));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_avgpool_nhwc_bwd(const float* dy, float* dx, int N, int S, int C, void* stream) {
  const int64_t total = (int64_t)N * S * C;
  DDL_CUDA_CHECK(ddl_launch_pdl(avgpool_bwd_kernel, dim3(grid_for(total, 256)), dim3(256), 0, (cudaStream_t)stream, dy, dx, total, S, C));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_im2col_stem7(const float* x, float* col, int N, int Cin, int H, int W, void* stream) {
  DDL_REQUIRE(Cin >= 1 && Cin <= 4, "im2col_stem7: Cin=%d outside [1,4]", Cin);
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;    // (H + 6 - 7) / 2 + 1
  DDL_CUDA_CHECK(ddl_launch_pdl(im2col_stem_kernel, dim3(grid_for((int64_t)N * Ho * Wo * 56, 256)), dim3(256), 0, (cudaStream_t)stream, x, reinterpret_cast<float4*>(col), N, Cin,
                                                                                             H, W, Ho, Wo));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
