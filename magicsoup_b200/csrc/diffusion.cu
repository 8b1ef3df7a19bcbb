// World.diffuse_molecules for sm_100a (reference python/magicsoup/world.py:627-665,
// weights from World._get_diffuse :948-972 and World._get_permeate :935-946), and the
// write-back half of World.enzymatic_activity (:624-625).
//
//   pass 1  stencil      shared-memory tiled circular 3x3 stencil of every molecule plane
//                        into scratch + exact fixed-point plane sums before/after
//   pass 2  correct      out += (sum_before - sum_after) / (H*W); clamp(0); back in place
//   pass 3  permeate     per cell and molecule: symmetric exchange with its pixel
//
// Spec order (bit-identical to oracle/ms_oracle.c): top=(nw+n)+ne, mid=w+e,
// bot=(sw+s)+se, nb=(top+mid)+bot, out=b*c+a*nb without FMA; plane sums are exact
// integers of 2^-32 (order independent, so atomics are deterministic), rounded once
// to float32 where the reference has torch.sum's float32 scalars (world.py:640-648).
#include "ms_common.cuh"

namespace ms {

constexpr int TW = 128;      // tile width  (32 lanes x float4)
constexpr int TH = 32;       // tile height (8 warps x 4 rows)
constexpr int TSTRIDE = 136; // smem row stride in floats; interior starts at column 4
constexpr int DTHREADS = 256;

struct FixSum {
  long long hi, lo;
  int bad;  // a value the fixed point cannot hold: NaN, Inf or |x| >= 2^31
  __device__ __forceinline__ void add(float x) {
    bad |= !(fabsf(x) < 2147483648.0f);
    const long long q = __double2ll_rn((double)x * 4294967296.0);
    hi += q >> 32;
    lo += q & 0xffffffffLL;
  }
};

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

__device__ __forceinline__ float stencil_point(float nw, float n, float ne, float wv, float c,
                                               float e, float sw, float sv, float se, float a,
                                               float b) {
  const float top = (nw + n) + ne;
  const float mid = wv + e;
  const float bot = (sw + sv) + se;
  const float nb = (top + mid) + bot;
  return b * c + a * nb;
}

// block-level reduction of the four fixed-point accumulators + one atomic each
__device__ void block_commit_sums(FixSum in, FixSum out, unsigned long long *sums /* 4 per plane */,
                                  unsigned long long *bad_flag /* 1 per plane */) {
  if (__syncthreads_or(in.bad | out.bad) && threadIdx.x == 0) atomicAdd(bad_flag, 1ull);
  __shared__ long long red[4][DTHREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long v0 = warp_sum_ll(in.hi), v1 = warp_sum_ll(in.lo), v2 = warp_sum_ll(out.hi),
            v3 = warp_sum_ll(out.lo);
  if (lane == 0) { red[0][warp] = v0; red[1][warp] = v1; red[2][warp] = v2; red[3][warp] = v3; }
  __syncthreads();
  if (threadIdx.x < 4) {
    long long t = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += red[threadIdx.x][w];
    atomicAdd(&sums[threadIdx.x], (unsigned long long)t);
  }
}

// Row `gx` of plane k of a slab: rows -1 and H come from the neighbours' halo rows
// (m,W) when given (multi-GPU row slabs), else they wrap around inside the plane.
__device__ __forceinline__ const float *slab_row(const float *plane, const float *halo_top,
                                                 const float *halo_bot, int k, int gx, int H, int W) {
  if (gx < 0) return halo_top ? halo_top + (size_t)k * W : plane + (size_t)(H - 1) * W;
  if (gx >= H) return halo_bot ? halo_bot + (size_t)k * W : plane;
  return plane + (size_t)gx * W;
}

// fast path: W % 128 == 0, H % 32 == 0.  grid = (W/TW, H/TH, m)
__global__ void __launch_bounds__(DTHREADS) stencil_tiled_kernel(const float *__restrict__ map,
                                                                 const float *__restrict__ halo_top,
                                                                 const float *__restrict__ halo_bot,
                                                                 float *__restrict__ out,
                                                                 const float *__restrict__ wa,
                                                                 const float *__restrict__ wb, int H,
                                                                 int W, unsigned long long *sums,
                                                                 const float *__restrict__ decay) {
  __shared__ __align__(16) float tile[(TH + 2) * TSTRIDE];
  const int k = blockIdx.z;
  const float dk = decay ? decay[k] : 1.0f;  // World.degrade_molecules folded in (x * 1 is exact)
  const int x0 = blockIdx.y * TH, y0 = blockIdx.x * TW;
  const float *plane = map + (size_t)k * H * W;
  float *oplane = out + (size_t)k * H * W;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  // interior columns of rows -1..TH: one float4 per lane per row
  for (int r = warp; r < TH + 2; r += DTHREADS / 32) {
    const float *src = slab_row(plane, halo_top, halo_bot, k, x0 + r - 1, H, W);
    float4 v = *reinterpret_cast<const float4 *>(src + y0 + 4 * lane);
    v.x = v.x * dk; v.y = v.y * dk; v.z = v.z * dk; v.w = v.w * dk;
    *reinterpret_cast<float4 *>(&tile[r * TSTRIDE + 4 + 4 * lane]) = v;
  }
  // halo columns -1 and TW
  for (int t = threadIdx.x; t < 2 * (TH + 2); t += DTHREADS) {
    const int r = t >> 1, side = t & 1;
    const float *src = slab_row(plane, halo_top, halo_bot, k, x0 + r - 1, H, W);
    int gy = side ? y0 + TW : y0 - 1;
    gy = gy < 0 ? gy + W : (gy >= W ? gy - W : gy);
    tile[r * TSTRIDE + (side ? 4 + TW : 3)] = src[gy] * dk;
  }
  __syncthreads();

  const float a = wa[k], b = wb[k];
  FixSum sin = {0, 0, 0}, sout = {0, 0, 0};
  const int c0 = 4 + 4 * lane;
  // rolling 3-row window: rows r-1, r, r+1 of the tile; each row = left, 4 centre, right
  float prev[6], cur[6], nxt[6];
  auto load_row = [&](int r, float *dst) {
    const float4 v = *reinterpret_cast<const float4 *>(&tile[r * TSTRIDE + c0]);
    dst[0] = tile[r * TSTRIDE + c0 - 1];
    dst[1] = v.x; dst[2] = v.y; dst[3] = v.z; dst[4] = v.w;
    dst[5] = tile[r * TSTRIDE + c0 + 4];
  };
  const int r0 = 4 * warp;  // first output row of this warp (tile row index r0+1)
  load_row(r0, prev);
  load_row(r0 + 1, cur);
#pragma unroll
  for (int rr = 0; rr < 4; rr++) {
    load_row(r0 + rr + 2, nxt);
    float o[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
      o[q] = stencil_point(prev[q], prev[q + 1], prev[q + 2], cur[q], cur[q + 1], cur[q + 2], nxt[q],
                           nxt[q + 1], nxt[q + 2], a, b);
      sin.add(cur[q + 1]);
      sout.add(o[q]);
    }
    *reinterpret_cast<float4 *>(oplane + (size_t)(x0 + r0 + rr) * W + y0 + 4 * lane) =
        make_float4(o[0], o[1], o[2], o[3]);
#pragma unroll
    for (int q = 0; q < 6; q++) { prev[q] = cur[q]; cur[q] = nxt[q]; }
  }
  block_commit_sums(sin, sout, sums + 4 * k, sums + 4 * gridDim.z + k);
}

// generic path: any H, W (e.g. the 5x5 map of the reference's test).  grid = (blocks, 1, m)
__global__ void __launch_bounds__(DTHREADS) stencil_generic_kernel(const float *__restrict__ map,
                                                                   const float *__restrict__ halo_top,
                                                                   const float *__restrict__ halo_bot,
                                                                   float *__restrict__ out,
                                                                   const float *__restrict__ wa,
                                                                   const float *__restrict__ wb, int H,
                                                                   int W, unsigned long long *sums,
                                                                   const float *__restrict__ decay) {
  const int k = blockIdx.z;
  const float *plane = map + (size_t)k * H * W;
  float *oplane = out + (size_t)k * H * W;
  const float a = wa[k], b = wb[k];
  const float dk = decay ? decay[k] : 1.0f;
  FixSum sin = {0, 0, 0}, sout = {0, 0, 0};
  const long long n = (long long)H * W;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n;
       t += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(t / W), y = (int)(t % W);
    const int ym = y == 0 ? W - 1 : y - 1, yp = y == W - 1 ? 0 : y + 1;
    const float *rm = slab_row(plane, halo_top, halo_bot, k, x - 1, H, W);
    const float *r0 = plane + (size_t)x * W;
    const float *rp = slab_row(plane, halo_top, halo_bot, k, x + 1, H, W);
    const float c = r0[y] * dk;
    const float o = stencil_point(rm[ym] * dk, rm[y] * dk, rm[yp] * dk, r0[ym] * dk, c, r0[yp] * dk, rp[ym] * dk,
                                  rp[y] * dk, rp[yp] * dk, a, b);
    oplane[t] = o;
    sin.add(c);
    sout.add(o);
  }
  block_commit_sums(sin, sout, sums + 4 * k, sums + 4 * gridDim.z + k);
}

// pass 2: world.py:648-649.  grid = (blocks, 1, m)
__global__ void __launch_bounds__(DTHREADS) correct_kernel(const float *__restrict__ out,
                                                           float *__restrict__ map, long long n_pxls,
                                                           long long n_pxls_total,
                                                           const unsigned long long *sums, int vec,
                                                           const unsigned *gate) {
  if (gate_closed(gate)) return;
  const int k = blockIdx.z;
  const long long *sk = reinterpret_cast<const long long *>(sums) + 4 * k;
  const float s0 = (float)((double)sk[0] + (double)sk[1] * 2.3283064365386963e-10);
  const float s1 = (float)((double)sk[2] + (double)sk[3] * 2.3283064365386963e-10);
  float corr = (s0 - s1) / (float)n_pxls_total;
  // a NaN / Inf in the plane makes the reference's float32 torch.sum NaN (world.py:640-648): whole plane NaN
  if (reinterpret_cast<const long long *>(sums)[4 * gridDim.z + k] != 0) corr = __int_as_float(0x7fc00000);
  const float *src = out + (size_t)k * n_pxls;
  float *dst = map + (size_t)k * n_pxls;
  const long long stride = (long long)gridDim.x * blockDim.x;
  if (vec) {  // n_pxls % 4 == 0 and both planes 16-byte aligned (correct_launch)
    const float4 *s4 = reinterpret_cast<const float4 *>(src);
    float4 *d4 = reinterpret_cast<float4 *>(dst);
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < (n_pxls >> 2); t += stride) {
      float4 v = s4[t];
      v.x = clamp0(v.x + corr); v.y = clamp0(v.y + corr);
      v.z = clamp0(v.z + corr); v.w = clamp0(v.w + corr);
      d4[t] = v;
    }
  } else {
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n_pxls; t += stride)
      dst[t] = clamp0(src[t] + corr);
  }
}

// pass 3: world.py:654-665, one thread per (cell, molecule)
__global__ void permeate_kernel(const float *cm_in, float *cm_out, float *map, const int *pos,
                                const float *perm, long long c, int m, int H, int W, const float *decay,
                                const unsigned *gate) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= c * m || gate_closed(gate)) return;
  const long long cell = r / m;
  const int k = (int)(r % m);
  const int x = pos[2 * cell], y = pos[2 * cell + 1];
  if ((unsigned)x >= (unsigned)H || (unsigned)y >= (unsigned)W) {  // outside the map: no exchange, no stray write
    cm_out[r] = decay ? cm_in[r] * decay[k] : cm_in[r];
    return;
  }
  const size_t mi = ((size_t)k * H + x) * W + y;
  const float xi = decay ? cm_in[r] * decay[k] : cm_in[r], xe = map[mi], q = perm[k];
  const float di = xi * q, de = xe * q;
  cm_out[r] = xi + (de - di);
  map[mi] = xe + (di - de);
}

// World.degrade_molecules, world.py:667-678: map[k] *= f[k], cell_molecules[:, k] *= f[k]
__global__ void degrade_kernel(float *map, long long n_pxls, float *cm, long long c, int m, const float *f,
                               const unsigned *gate) {
  if (gate_closed(gate)) return;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (int k = 0; k < m; k++) {
    const float fk = f[k];
    float *plane = map + (size_t)k * n_pxls;
    for (long long t = t0; t < n_pxls; t += stride) plane[t] = plane[t] * fk;
  }
  for (long long t = t0; t < c * m; t += stride) cm[t] = cm[t] * f[t % m];
}

// ---------------------------------------------------------------------------
// Pixel-ordered variants for densely populated maps.  The kernels above take one thread per (cell,
// molecule): neighbouring threads touch planes H*W apart, i.e. one 4-byte value per 32-byte sector
// on the map side.  With a pixel -> cell index (ms_build_pixel_index) one thread takes one PIXEL and
// walks the m planes, so every plane access of a warp is one coalesced row segment and the cell side
// is the cell's whole m-float row.  Same arithmetic per (cell, molecule), so the same bits.  m <= 32.
// ---------------------------------------------------------------------------
__global__ void pixel_index_kernel(const int *pos, long long c, int H, int W, int *pixel_cell) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= c) return;
  const int x = pos[2 * r], y = pos[2 * r + 1];
  if ((unsigned)x < (unsigned)H && (unsigned)y < (unsigned)W) pixel_cell[(size_t)x * W + y] = (int)r;  // else: no pixel
}

// A warp takes 32 consecutive pixels.  Map side: lane L owns pixel L, one coalesced 128-byte
// access per plane.  Cell side: the 32 lanes read / write ONE cell's row together (coalesced), pixel
// after pixel.  A (32 x 33)-float tile in shared memory transposes between the two.  m <= 32.
constexpr int PIX_WARPS = 8;

// lanes are split into groups of G = 2^ceil(log2(n)) lanes (n values per cell); a warp handles 32 / G cells
// per pass, so that no lane idles when a cell's row is shorter than the warp
__device__ __forceinline__ int lane_group_size(int n) {
  int g = 1;
  while (g < n) g <<= 1;
  return g < 32 ? g : 32;
}

__global__ void __launch_bounds__(32 * PIX_WARPS) permeate_pixel_kernel(const float *__restrict__ cm_in,
                                                                        float *__restrict__ cm_out, float *map,
                                                                        const int *__restrict__ pixel_cell,
                                                                        const float *__restrict__ perm, int m,
                                                                        long long n_pxls, const float *__restrict__ decay,
                                                                        const unsigned *gate) {
  __shared__ float tile[PIX_WARPS][32][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long px0 = ((long long)blockIdx.x * PIX_WARPS + warp) * 32;
  if (px0 >= n_pxls || gate_closed(gate)) return;
  const long long px = px0 + lane;
  const bool inside = px < n_pxls;
  const int cell = inside ? pixel_cell[px] : -1;
  if (__ballot_sync(0xffffffffu, cell >= 0) == 0u) return;
  float (*t)[33] = tile[warp];
#pragma unroll 4
  for (int k = 0; k < m; k++) t[lane][k] = inside ? map[(size_t)k * n_pxls + px] : 0.0f;
  __syncwarp();
  const int G = lane_group_size(m), li = lane & (G - 1), grp = lane / G, per_pass = 32 / G;
  const float q = li < m ? perm[li] : 0.0f, dk = (decay && li < m) ? decay[li] : 1.0f;
#pragma unroll 4
  for (int j0 = 0; j0 < 32; j0 += per_pass) {
    const int j = j0 + grp;  // the pixel (of this warp's 32) whose cell this lane group exchanges with
    const int cj = __shfl_sync(0xffffffffu, cell, j);
    if (cj < 0 || li >= m) continue;
    const float raw = cm_in[(size_t)cj * m + li];
    const float xi = decay ? raw * dk : raw, xe = t[j][li];
    const float di = xi * q, de = xe * q;
    cm_out[(size_t)cj * m + li] = xi + (de - di);
    t[j][li] = xe + (di - de);
  }
  __syncwarp();
  if (inside && cell >= 0) {
#pragma unroll 4
    for (int k = 0; k < m; k++) map[(size_t)k * n_pxls + px] = t[lane][k];
  }
}

__global__ void __launch_bounds__(32 * PIX_WARPS) scatter_pixel_kernel(const float *__restrict__ X,
                                                                       const int *__restrict__ pixel_cell, int m,
                                                                       long long n_pxls, float *__restrict__ cm_out,
                                                                       float *__restrict__ map, const unsigned *gate) {
  __shared__ float tile[PIX_WARPS][32][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long px0 = ((long long)blockIdx.x * PIX_WARPS + warp) * 32;
  if (px0 >= n_pxls || gate_closed(gate)) return;
  const long long px = px0 + lane;
  const int cell = px < n_pxls ? pixel_cell[px] : -1;
  if (__ballot_sync(0xffffffffu, cell >= 0) == 0u) return;
  float (*t)[33] = tile[warp];
  const int s = 2 * m;
  if (s <= 32) {  // a lane group per cell: 32 / G cells per pass
    const int G = lane_group_size(s), li = lane & (G - 1), grp = lane / G, per_pass = 32 / G;
#pragma unroll 4
    for (int j0 = 0; j0 < 32; j0 += per_pass) {
      const int j = j0 + grp;
      const int cj = __shfl_sync(0xffffffffu, cell, j);
      if (cj < 0 || li >= s) continue;
      const float v = X[(size_t)cj * s + li];
      if (li < m) cm_out[(size_t)cj * m + li] = v;
      else t[j][li - m] = v;
    }
  } else {
    for (int j = 0; j < 32; j++) {
      const int cj = __shfl_sync(0xffffffffu, cell, j);
      if (cj < 0) continue;
      const float *x = X + (size_t)cj * s;
      for (int i = lane; i < s; i += 32) {
        const float v = x[i];
        if (i < m) cm_out[(size_t)cj * m + i] = v;
        else t[j][i - m] = v;
      }
    }
  }
  __syncwarp();
  if (cell >= 0) {
#pragma unroll 4
    for (int k = 0; k < m; k++) map[(size_t)k * n_pxls + px] = t[lane][k];
  }
}

// world.py:624-625, one thread per (cell, signal)
__global__ void scatter_kernel(const float *X, const int *pos, long long c, int m, int H, int W,
                               float *cm_out, float *map, const unsigned *gate) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int s = 2 * m;
  if (r >= c * s || gate_closed(gate)) return;
  const long long cell = r / s;
  const int i = (int)(r % s);
  const float v = X[r];
  if (i < m) {
    cm_out[cell * m + i] = v;
  } else {
    const int x = pos[2 * cell], y = pos[2 * cell + 1];
    if ((unsigned)x < (unsigned)H && (unsigned)y < (unsigned)W) map[((size_t)(i - m) * H + x) * W + y] = v;
  }
}

}  // namespace ms

using namespace ms;

#include <stdio.h>
static thread_local char g_static_too_large_msg[320] = "";

// Step gate (ms_step_gate): while one is set for the calling thread, the kernels that COMMIT results
// (write-back, correction, permeation, degradation) are launched with it and leave without writing
// when the word it points to is 0 at the time they run.  It lets a host enqueue the rest of a step
// behind the kinetics launch before it has read the early-exit flags (ms_early_exit_verdict).
static thread_local const unsigned *g_step_gate = nullptr;
extern "C" void ms_step_gate(const uint32_t *gate) { g_step_gate = gate; }

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// m * H * W without wrapping: false if the product does not fit `limit`
static bool volume_fits(int64_t m, int64_t H, int64_t W, int64_t limit, int64_t *volume) {
  if (m < 0 || H <= 0 || W <= 0 || H > limit / W) return false;
  const int64_t plane = H * W;
  if (m > 0 && plane > limit / m) return false;
  *volume = m * plane;
  return true;
}

extern "C" size_t ms_diffuse_scratch_bytes(int64_t m, int64_t H, int64_t W) {
  if (m <= 0 || H <= 0 || W <= 0) return 256;
  int64_t vol = 0;
  if (!volume_fits(m, H, W, (int64_t)1 << 58, &vol)) return SIZE_MAX;  // no buffer is ever large enough
  return align256((size_t)vol * sizeof(float)) + align256((size_t)m * 5 * sizeof(long long));
}

static inline bool aligned16(const void *p) { return ((uintptr_t)p & 15) == 0; }

static int stencil_launch(const float *map, const float *halo_top, const float *halo_bot,
                          const float *a, const float *b, int64_t m, int64_t H, int64_t W, float *tmp,
                          unsigned long long *sums, cudaStream_t st, const float *decay = nullptr) {
  MS_CUDA_CHECK(cudaMemsetAsync(sums, 0, (size_t)m * 5 * sizeof(long long), st));
  const long long n = (long long)H * W;
  // the tiled kernel moves float4: rows must start 16-byte aligned in every buffer it touches
  const bool vec_ok = aligned16(map) && aligned16(tmp) && (!halo_top || aligned16(halo_top)) &&
                      (!halo_bot || aligned16(halo_bot));
  if (W % TW == 0 && H % TH == 0 && vec_ok) {
    dim3 grid((unsigned)(W / TW), (unsigned)(H / TH), (unsigned)m);
    stencil_tiled_kernel<<<grid, DTHREADS, 0, st>>>(map, halo_top, halo_bot, tmp, a, b, (int)H, (int)W, sums, decay);
  } else {
    unsigned blocks = (unsigned)((n + DTHREADS - 1) / DTHREADS);
    if (blocks > 1024) blocks = 1024;
    dim3 grid(blocks, 1, (unsigned)m);
    stencil_generic_kernel<<<grid, DTHREADS, 0, st>>>(map, halo_top, halo_bot, tmp, a, b, (int)H, (int)W, sums, decay);
  }
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

static int correct_launch(const float *tmp, float *map, int64_t m, int64_t H, int64_t W,
                          long long n_pxls_total, const unsigned long long *sums, cudaStream_t st,
                          const unsigned *gate) {
  const long long n = (long long)H * W;
  const bool vec = (n & 3) == 0 && aligned16(tmp) && aligned16(map);
  long long work = vec ? n / 4 : n;
  unsigned blocks = (unsigned)((work + DTHREADS - 1) / DTHREADS);
  if (blocks > 148 * 8) blocks = 148 * 8;
  correct_kernel<<<dim3(blocks, 1, (unsigned)m), DTHREADS, 0, st>>>(tmp, map, n, n_pxls_total, sums, vec ? 1 : 0, gate);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

static bool bad_map(int64_t m, int64_t H, int64_t W) {
  return m < 0 || H <= 0 || W <= 0 || H > INT32_MAX || W > INT32_MAX || m > 65535;
}

extern "C" int ms_diffuse_molecules(float *molecule_map, const float *a, const float *b, int64_t m,
                                    int64_t H, int64_t W, void *scratch, size_t scratch_bytes,
                                    void *stream) {
  if (bad_map(m, H, W)) return MS_ERR_INVALID_ARGUMENT;
  if (m == 0) return MS_OK;
  if (!molecule_map || !a || !b) return MS_ERR_INVALID_ARGUMENT;
  if (!scratch || scratch_bytes < ms_diffuse_scratch_bytes(m, H, W)) return MS_ERR_SCRATCH_TOO_SMALL;
  float *tmp = (float *)scratch;
  unsigned long long *sums =
      (unsigned long long *)((char *)scratch + align256((size_t)m * H * W * sizeof(float)));
  int rc = stencil_launch(molecule_map, nullptr, nullptr, a, b, m, H, W, tmp, sums, (cudaStream_t)stream);
  if (rc != MS_OK) return rc;
  return correct_launch(tmp, molecule_map, m, H, W, (long long)H * W, sums, (cudaStream_t)stream, g_step_gate);
}

static int diffuse_stencil_impl(const float *map_slab, const float *halo_top, const float *halo_bottom,
                                const float *a, const float *b, const float *decay, int64_t m, int64_t H,
                                int64_t W, float *out, int64_t *sums, void *stream) {
  if (bad_map(m, H, W)) return MS_ERR_INVALID_ARGUMENT;
  if (m == 0) return MS_OK;
  if (!map_slab || !a || !b || !out || !sums || out == map_slab) return MS_ERR_INVALID_ARGUMENT;
  return stencil_launch(map_slab, halo_top, halo_bottom, a, b, m, H, W, out, (unsigned long long *)sums,
                        (cudaStream_t)stream, decay);
}

extern "C" int ms_diffuse_stencil(const float *map_slab, const float *halo_top, const float *halo_bottom,
                                  const float *a, const float *b, int64_t m, int64_t H, int64_t W,
                                  float *out, int64_t *sums, void *stream) {
  return diffuse_stencil_impl(map_slab, halo_top, halo_bottom, a, b, nullptr, m, H, W, out, sums, stream);
}

extern "C" int ms_degrade_diffuse_stencil(const float *map_slab, const float *halo_top, const float *halo_bottom,
                                          const float *a, const float *b, const float *decay, int64_t m,
                                          int64_t H, int64_t W, float *out, int64_t *sums, void *stream) {
  if (!decay && m > 0) return MS_ERR_INVALID_ARGUMENT;
  return diffuse_stencil_impl(map_slab, halo_top, halo_bottom, a, b, decay, m, H, W, out, sums, stream);
}

extern "C" int ms_diffuse_correct(const float *out, float *map_slab, const int64_t *sums, int64_t m,
                                  int64_t H, int64_t W, int64_t n_pxls_total, void *stream) {
  if (bad_map(m, H, W) || n_pxls_total <= 0) return MS_ERR_INVALID_ARGUMENT;
  if (m == 0) return MS_OK;
  if (!out || !map_slab || !sums) return MS_ERR_INVALID_ARGUMENT;
  return correct_launch(out, map_slab, m, H, W, n_pxls_total, (const unsigned long long *)sums,
                        (cudaStream_t)stream, g_step_gate);
}

static int permeate_impl(const float *cell_molecules_in, float *cell_molecules_out, float *molecule_map,
                         const int32_t *cell_positions, const float *perm, const float *decay, int64_t c, int64_t m,
                         int64_t H, int64_t W, void *stream) {
  if (c < 0 || m < 0 || H <= 0 || W <= 0 || H > INT32_MAX || W > INT32_MAX) return MS_ERR_INVALID_ARGUMENT;
  if (c * m == 0) return MS_OK;
  if (!cell_molecules_in || !cell_molecules_out || !molecule_map || !cell_positions || !perm)
    return MS_ERR_INVALID_ARGUMENT;
  const long long n = c * m;
  permeate_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      cell_molecules_in, cell_molecules_out, molecule_map, cell_positions, perm, c, (int)m, (int)H, (int)W, decay,
      g_step_gate);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_permeate(const float *cell_molecules_in, float *cell_molecules_out,
                           float *molecule_map, const int32_t *cell_positions, const float *perm,
                           int64_t c, int64_t m, int64_t H, int64_t W, void *stream) {
  return permeate_impl(cell_molecules_in, cell_molecules_out, molecule_map, cell_positions, perm, nullptr, c, m, H, W,
                       stream);
}

extern "C" int ms_degrade_permeate(const float *cell_molecules_in, float *cell_molecules_out, float *molecule_map,
                                   const int32_t *cell_positions, const float *perm, const float *decay, int64_t c,
                                   int64_t m, int64_t H, int64_t W, void *stream) {
  if (!decay && c * m > 0) return MS_ERR_INVALID_ARGUMENT;
  return permeate_impl(cell_molecules_in, cell_molecules_out, molecule_map, cell_positions, perm, decay, c, m, H, W,
                       stream);
}

extern "C" int ms_build_pixel_index(const int32_t *cell_positions, int64_t c, int64_t H, int64_t W,
                                    int32_t *pixel_cell, void *stream) {
  if (c < 0 || H <= 0 || W <= 0 || H > INT32_MAX || W > INT32_MAX || c > INT32_MAX || !pixel_cell ||
      (c > 0 && !cell_positions))
    return MS_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  MS_CUDA_CHECK(cudaMemsetAsync(pixel_cell, 0xff, (size_t)H * W * sizeof(int32_t), st));  // -1: no cell
  if (c > 0) pixel_index_kernel<<<(unsigned)((c + 255) / 256), 256, 0, st>>>(cell_positions, c, (int)H, (int)W, pixel_cell);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_permeate_by_pixel(const float *cell_molecules_in, float *cell_molecules_out, float *molecule_map,
                                    const int32_t *pixel_cell, const float *perm, const float *decay, int64_t m,
                                    int64_t H, int64_t W, void *stream) {
  if (bad_map(m, H, W)) return MS_ERR_INVALID_ARGUMENT;
  if (m == 0) return MS_OK;
  if (!cell_molecules_in || !cell_molecules_out || !molecule_map || !pixel_cell || !perm)
    return MS_ERR_INVALID_ARGUMENT;
  if (m > 32) return MS_ERR_TOO_LARGE;  // the transposing tile holds 32 planes; use ms_permeate
  const long long n = (long long)H * W;
  permeate_pixel_kernel<<<(unsigned)((n + 32 * PIX_WARPS - 1) / (32 * PIX_WARPS)), 32 * PIX_WARPS, 0, (cudaStream_t)stream>>>(
      cell_molecules_in, cell_molecules_out, molecule_map, pixel_cell, perm, (int)m, n, decay, g_step_gate);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_scatter_signals_by_pixel(const float *X, const int32_t *pixel_cell, int64_t m, int64_t H,
                                           int64_t W, float *cell_molecules_out, float *molecule_map,
                                           void *stream) {
  if (bad_map(m, H, W)) return MS_ERR_INVALID_ARGUMENT;
  if (m == 0) return MS_OK;
  if (!X || !pixel_cell || !cell_molecules_out || !molecule_map) return MS_ERR_INVALID_ARGUMENT;
  if (m > 32) return MS_ERR_TOO_LARGE;  // the transposing tile holds 32 planes; use ms_scatter_signals
  const long long n = (long long)H * W;
  scatter_pixel_kernel<<<(unsigned)((n + 32 * PIX_WARPS - 1) / (32 * PIX_WARPS)), 32 * PIX_WARPS, 0,
                         (cudaStream_t)stream>>>(X, pixel_cell, (int)m, n, cell_molecules_out, molecule_map,
                                                 g_step_gate);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_degrade_molecules(float *molecule_map, float *cell_molecules, const float *decay, int64_t c,
                                    int64_t m, int64_t H, int64_t W, void *stream) {
  if (bad_map(m, H, W) || c < 0) return MS_ERR_INVALID_ARGUMENT;
  if (m == 0) return MS_OK;
  if (!molecule_map || !decay || (c > 0 && !cell_molecules)) return MS_ERR_INVALID_ARGUMENT;
  degrade_kernel<<<148 * 4, 256, 0, (cudaStream_t)stream>>>(molecule_map, (long long)H * W, cell_molecules, c, (int)m,
                                                           decay, g_step_gate);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_scatter_signals(const float *X, const int32_t *cell_positions, int64_t c, int64_t m,
                                  int64_t H, int64_t W, float *cell_molecules_out, float *molecule_map,
                                  void *stream) {
  if (c < 0 || m < 0 || H <= 0 || W <= 0 || H > INT32_MAX || W > INT32_MAX) return MS_ERR_INVALID_ARGUMENT;
  if (c * m == 0) return MS_OK;
  if (!X || !cell_positions || !cell_molecules_out || !molecule_map) return MS_ERR_INVALID_ARGUMENT;
  const long long n = c * 2 * m;
  scatter_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      X, cell_positions, c, (int)m, (int)H, (int)W, cell_molecules_out, molecule_map, g_step_gate);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

// ---------------------------------------------------------------------------
// Diffusion without any caller-provided or allocated scratch.
//
// Who needs this: hosts that embed the library but may not allocate device memory when they
// call it -- for example an application that hands all device memory to its own pool at
// start-up and forbids cudaMalloc afterwards, or a step loop that must not hit the implicit
// device synchronisation of cudaMalloc/cudaFree.  ms_diffuse_molecules makes such a caller
// carve ms_diffuse_scratch_bytes() out of its own memory; this variant needs nothing: the
// stencil output and the plane sums live in a fixed array inside the library image.
//
// The price: the array has a fixed size (MS_STATIC_DIFFUSE_MAX_FLOATS floats = 16 MiB, plus
// MS_STATIC_DIFFUSE_MAX_PLANES plane sums) that every GPU which loads the library holds
// until the process exits, and ALL calls on one GPU share it: two calls that overlap on the
// device (different streams) race on it.  The array keeps whatever the previous call left;
// the stencil output is fully rewritten before it is read, and the plane sums are cleared by
// every call before use.
// ---------------------------------------------------------------------------
#include <mutex>

#define MS_STATIC_DIFFUSE_MAX_FLOATS (4ll * 1024 * 1024)
#define MS_STATIC_DIFFUSE_MAX_PLANES 256

static __device__ __align__(256) float g_static_diffuse_out[MS_STATIC_DIFFUSE_MAX_FLOATS];
static __device__ __align__(256) unsigned long long g_static_diffuse_sums[5 * MS_STATIC_DIFFUSE_MAX_PLANES];

namespace {
constexpr int kMaxDevices = 64;
struct StaticScratch {
  float *out = nullptr;
  unsigned long long *sums = nullptr;
};
StaticScratch g_static_scratch[kMaxDevices];
std::mutex g_static_scratch_mutex;
}  // namespace

// The names above are host-side stand-ins in host code, not device addresses: the device
// address of the copy that lives on the CURRENT device is looked up once per device.
static int static_scratch_for_current_device(StaticScratch &res) {
  int dev = -1;
  MS_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return MS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (g_static_scratch[dev].out == nullptr) {
    void *out = nullptr, *sums = nullptr;
    MS_CUDA_CHECK(cudaGetSymbolAddress(&out, g_static_diffuse_out));
    MS_CUDA_CHECK(cudaGetSymbolAddress(&sums, g_static_diffuse_sums));
    g_static_scratch[dev].out = (float *)out;
    g_static_scratch[dev].sums = (unsigned long long *)sums;
  }
  res = g_static_scratch[dev];
  return MS_OK;
}

extern "C" int64_t ms_diffuse_static_max_floats(void) { return MS_STATIC_DIFFUSE_MAX_FLOATS; }

extern "C" int ms_diffuse_molecules_static(float *molecule_map, const float *a, const float *b,
                                           int64_t m, int64_t H, int64_t W, void *stream) {
  if (bad_map(m, H, W)) return MS_ERR_INVALID_ARGUMENT;
  if (m == 0) return MS_OK;
  if (!molecule_map || !a || !b) return MS_ERR_INVALID_ARGUMENT;
  g_static_too_large_msg[0] = 0;
  int64_t vol = 0;  // compared by division: the product of three int64 can wrap below the limit
  if (m > MS_STATIC_DIFFUSE_MAX_PLANES || !volume_fits(m, H, W, MS_STATIC_DIFFUSE_MAX_FLOATS, &vol)) {
    snprintf(g_static_too_large_msg, sizeof(g_static_too_large_msg),
             "ms_diffuse_molecules_static: map of %lld x %lld x %lld floats exceeds the built-in scratch of "
             "%lld floats / %d planes; use ms_diffuse_molecules with caller scratch",
             (long long)m, (long long)H, (long long)W, (long long)MS_STATIC_DIFFUSE_MAX_FLOATS,
             MS_STATIC_DIFFUSE_MAX_PLANES);
    return MS_ERR_TOO_LARGE;
  }
  StaticScratch sc;
  int rc = static_scratch_for_current_device(sc);
  if (rc != MS_OK) return rc;
  // stencil_launch clears the plane sums (cudaMemsetAsync) before accumulating into them
  rc = stencil_launch(molecule_map, nullptr, nullptr, a, b, m, H, W, sc.out, sc.sums, (cudaStream_t)stream);
  if (rc != MS_OK) return rc;
  return correct_launch(sc.out, molecule_map, m, H, W, (long long)H * W, sc.sums, (cudaStream_t)stream, nullptr);
}

extern "C" const char *ms_diffuse_static_last_error(void) { return g_static_too_large_msg; }
