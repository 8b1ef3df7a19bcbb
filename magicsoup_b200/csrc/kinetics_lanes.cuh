// Lane path of the fused kinetics kernel (included by kinetics.cu): one warp integrates one cell
// WITHOUT any CTA barrier, on fixed-format lists that make every lane's work uniform:
//
//   T     power table of the current signals: T[0] = 1, T[1 + (n-1)*S2 + i] = X_i^n for n = 1..3,
//         then one slot per (signal, exponent >= 4) entry of the cell ("rare" slots)
//   ent8  per ACTIVE protein t (ascending protein index): 4 + 4 byte indices into T, substrate
//         side and product side in ascending signal order, padded with index 0 (x * 1 is exact);
//         proteins with more than 4 entries on a side continue in `ext8` chunks
//   csc   per signal: chunks of 4 (active protein t, N) entries in ascending protein order, 16 bit
//         each, padded with (dummy protein PA, N = 0); `nv` holds the 4 rates of each chunk
//   per-protein scalars Ke, Kmf, Kmb, Vmax gathered in active order, so lane t reads element t
//
// A product over a protein's entries is 8 table look-ups and 8 multiplications without a branch,
// a sum over a signal's entries is one 64-bit and one 128-bit load per 4 entries.  The arithmetic
// and its order are those of run_sparse (and of oracle/ms_oracle.c): products and sums run over the
// same entries in the same order, padding entries are exact no-ops.
//
// The static arrays (col .. csc) are what the list cache keeps per parameter row.
#pragma once

#ifndef MS_PRODUCTS_INLINE
#define MS_PRODUCTS_INLINE __forceinline__
#endif

namespace ms {

// header of a lane-path record / of the warp's current cell
struct LaneCounts {
  int n_act, n_rare, n_reg;
};
__device__ __forceinline__ unsigned pack_counts(const LaneCounts &c) {
  return (unsigned)c.n_act | ((unsigned)c.n_rare << 8) | ((unsigned)c.n_reg << 16);
}
__device__ __forceinline__ LaneCounts unpack_counts(unsigned w) {
  LaneCounts c;
  c.n_act = (int)(w & 0xffu);
  c.n_rare = (int)((w >> 8) & 0xffu);
  c.n_reg = (int)((w >> 16) & 0xffu);
  return c;
}

__device__ __forceinline__ void red_add_shared(unsigned addr, unsigned v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_u8(unsigned addr, unsigned v) {
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_u16r(unsigned addr, unsigned v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ uint2 lds_v2(unsigned addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ float4 lds_f4(unsigned addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f4(unsigned addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// ---------------------------------------------------------------------------
// Scan of the lane path: the cell's dense (p,s) tiles of Nf, Nb, N, A are streamed once, 512 bytes
// per tensor and step, staged through shared memory two steps ahead (cp.async, 16 bytes per lane,
// two slot sets of 4 x 512 bytes at `stage`), and every element with a non-zero value is pushed onto
// the queue (2 words per element, row-major order).  Half of the steps of a typical tile are all
// zero: such a step costs four 128-bit loads, an OR tree and one vote.  Returns the queue length or
// -1 (the tile does not fit the queue or the packing).
// ---------------------------------------------------------------------------
__device__ __noinline__ int scan_tile_lanes(const int *gNf, const int *gNb, const int *gN, const int *gA, int p, int s,
                                            unsigned stage, unsigned aQueue, int cap, int lane) {
  const int total = p * s;
  const unsigned lt = (1u << lane) - 1u;
  const unsigned set0 = stage + 16u * (unsigned)lane;
  const int drow = 128 / s, dsig = 128 - drow * s;
  int q0 = 4 * lane;  // this lane's 4 consecutive elements of the current step (s % 4 == 0: one row)
  int row = q0 / s, sig0 = q0 - row * s;
  int nq = 0, chk = 0;
#pragma unroll
  for (int g = 0; g < 2; g++) {  // two steps in flight
    const int q = q0 + 128 * g;
    const unsigned slots = set0 + 2048u * (unsigned)g;
    if (q < total) {
      cp_async16(slots, gNf + q, 0u);
      cp_async16(slots + 512u, gNb + q, 0u);
      cp_async16(slots + 1024u, gN + q, 0u);
      cp_async16(slots + 1536u, gA + q, 0u);
    }
    cp_async_commit();
  }
  unsigned toggle = 0;
#pragma unroll 1
  for (int base = 0; base < total; base += 128) {
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    const unsigned slots = set0 + toggle;
    toggle ^= 2048u;
    int nf[4], nb[4], nn[4], na[4];
    lds_vec4(slots, nf);
    lds_vec4(slots + 512u, nb);
    lds_vec4(slots + 1024u, nn);
    lds_vec4(slots + 1536u, na);
    const bool valid = q0 < total;  // past the end the slots hold stale words
    const int o0 = nf[0] | nb[0] | nn[0] | na[0], o1 = nf[1] | nb[1] | nn[1] | na[1];
    const int o2 = nf[2] | nb[2] | nn[2] | na[2], o3 = nf[3] | nb[3] | nn[3] | na[3];
    const unsigned any_l = valid ? (unsigned)(o0 | o1 | o2 | o3) : 0u;
    {  // refill the slots just read with the step after next (ordered behind the reads by `any_l`)
      const int q = q0 + 256;
      if (q < total) {
        cp_async16(slots, gNf + q, any_l);
        cp_async16(slots + 512u, gNb + q, any_l);
        cp_async16(slots + 1024u, gN + q, any_l);
        cp_async16(slots + 1536u, gA + q, any_l);
      }
      cp_async_commit();
    }
    const unsigned b1 = __ballot_sync(FULL, any_l != 0u);
    if (b1) {
      unsigned mm = (o0 != 0 ? 1u : 0u) | (o1 != 0 ? 2u : 0u) | (o2 != 0 ? 4u : 0u) | (o3 != 0 ? 8u : 0u);
      if (!valid) mm = 0u;
      const int cnt = __popc(mm);
      // exclusive prefix of cnt over the lanes: most lanes hold no element, few hold more than one
      int pos = nq + __popc(b1 & lt);
      nq += __popc(b1);
      const unsigned b2 = __ballot_sync(FULL, cnt >= 2);
      if (b2) {
        const unsigned b3 = __ballot_sync(FULL, cnt >= 3), b4 = __ballot_sync(FULL, cnt >= 4);
        pos += __popc(b2 & lt) + __popc(b3 & lt) + __popc(b4 & lt);
        nq += __popc(b2) + __popc(b3) + __popc(b4);
      }
      // a lane's elements one at a time (usually one): selected with a few SEL, packed, stored
#pragma unroll 1
      while (mm) {
        const int e = __ffs(mm) - 1;
        mm &= mm - 1u;
        const int vf = pick<4>(nf, e), vb = pick<4>(nb, e), vn = pick<4>(nn, e), va = pick<4>(na, e);
        // packing: 0 <= Nf, Nb < 256, |N| < 2^15, |A| < 2^7
        chk |= ((vf | vb) & 0xffffff00) | ((vn + 32768) & 0xffff0000) | ((va + 128) & 0xffffff00);
        if (pos < cap)
          queue_store(aQueue, pos, (unsigned)(sig0 + e) | ((unsigned)row << 8) | ((unsigned)(va & 0xff) << 16),
                      ((unsigned)vf & 0xffu) | (((unsigned)vb & 0xffu) << 8) | ((unsigned)vn << 16));
        pos++;
      }
    }
    q0 += 128;
    row += drow;
    sig0 += dsig;
    if (sig0 >= s) { sig0 -= s; row++; }
  }
  if (__any_sync(FULL, chk != 0) || nq > cap) return -1;
  return nq;
}

// ---------------------------------------------------------------------------
// Builds the fixed-format lists of one cell from the queue of its non-zero tile elements (scan
// phase 1: row-major order, i.e. ascending protein, then ascending signal).  Every position is a
// deterministic rank (ballot / match), never the return value of an atomic.  Returns false if the
// cell does not fit the format (it is then integrated by the tier-2 launch or the dense path).
// ---------------------------------------------------------------------------
__device__ bool build_lane_lists(const KinArgs &a, const LaneLayout &L, unsigned *b, unsigned sb, long long prm_row,
                                 int lane, int nq, LaneCounts &cnt) {
  const int p = a.p, s = a.s;
  const unsigned lt = (1u << lane) - 1u;
  (void)b;
  const unsigned aQueue = sb + 4u * (unsigned)L.oQueue;
  // temporaries in working arrays that are idle until the cell is integrated (shared-window addresses)
  const unsigned rowcnt = sb + 4u * (unsigned)L.oX0;        // [PA]  substrate | product << 8 | regulatory << 16 entries of a row
  const unsigned rowmap = rowcnt + 4u * (unsigned)L.PA;     // [PA]  active index of a row (0xff: inactive) | chunk/reg starts
  const unsigned rowrun = rowmap + 4u * (unsigned)L.PA;     // [PA]  entries placed so far: substrate | product << 8
  const unsigned sigcnt = rowrun + 4u * (unsigned)L.PA;     // [S2]  N entries of a signal, then its first chunk
  const unsigned sigrun = sigcnt + 4u * (unsigned)L.S2;     // [S2]  N entries placed so far
  // (3*PA + 2*S2 words from oX0 on: working arrays that are initialised after the build; the queue starts
  // behind them, make_lane_layout)
#pragma unroll 1
  for (int j = lane; j < 3 * L.PA + 2 * L.S2; j += 32) sts_u(rowcnt + 4u * (unsigned)j, 0u);
#pragma unroll 1
  for (int e = lane; e < 2 * L.PA + 2 * L.EX; e += 32) sts_u(sb + 4u * (unsigned)(L.oEnt8 + e), 0u);  // ent8, ext8
  {
    const unsigned pad = (unsigned)L.PA | ((unsigned)L.PA << 16);  // (dummy protein, N = 0) twice
#pragma unroll 1
    for (int e = lane; e < 2 * L.CH; e += 32) sts_u(sb + 4u * (unsigned)(L.oCsc + e), pad);
  }
  __syncwarp();
  // ---- pass A: counts
  int bad = 0, n_rare = 0;
#pragma unroll 1
  for (int base = 0; base < nq; base += 32) {
    const int k = base + lane;
    const bool v = k < nq;
    const uint2 qw = v ? lds_v2(aQueue + 8u * (unsigned)k) : make_uint2(0u, 0u);
    const unsigned sig = qw.x & 0xffu, row = (qw.x >> 8) & 0xffu;
    const int av = (int)(int8_t)((qw.x >> 16) & 0xffu);
    const unsigned nf = qw.y & 0xffu, nb = (qw.y >> 8) & 0xffu;
    const int n = (int)qw.y >> 16;
    const unsigned inc = (nf ? 1u : 0u) | (nb ? 1u << 8 : 0u) | (av ? 1u << 16 : 0u);
    if (inc) red_add_shared(rowcnt + 4u * row, inc);
    if (n != 0) red_add_shared(sigcnt + 4u * sig, 1u);
    bad |= (n < -256 || n > 255) ? 1 : 0;
    n_rare += __popc(__ballot_sync(FULL, nf > 3u)) + __popc(__ballot_sync(FULL, nb > 3u));
  }
  if (__any_sync(FULL, bad != 0) || n_rare > L.R) return false;
  __syncwarp();
  // ---- active proteins, their extension chunks and regulatory entries (prefix sums in row order)
  int n_act = 0, ext_tot = 0, reg_tot = 0;
#pragma unroll 1
  for (int j0 = 0; j0 < L.PA; j0 += 32) {
    const int j = j0 + lane;
    const unsigned cw = j < p ? lds_u(rowcnt + 4u * (unsigned)j) : 0u;
    const unsigned ns = cw & 0xffu, np = (cw >> 8) & 0xffu;
    const bool active = (ns | np) != 0u;
    const unsigned nr = active ? (cw >> 16) & 0xffu : 0u;
    const unsigned longest = ns > np ? ns : np;
    const unsigned n_ext = active ? (longest + 3u) / 4u - 1u : 0u;
    bad |= (n_ext > (unsigned)LANE_EXT_MAX || nr > (unsigned)LANE_REG_MAX) ? 1 : 0;
    const unsigned ab = __ballot_sync(FULL, active);
    unsigned incl = n_ext | (nr << 16);
    const unsigned own = incl;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned y = __shfl_up_sync(FULL, incl, d);
      if (lane >= d) incl += y;
    }
    const unsigned excl = incl - own;
    const unsigned t = (unsigned)n_act + __popc(ab & lt);
    const unsigned ext_start = (unsigned)ext_tot + (excl & 0xffffu), reg_start = (unsigned)reg_tot + (excl >> 16);
    if (j < L.PA) sts_u(rowmap + 4u * (unsigned)j, active ? (t | (ext_start << 8) | (reg_start << 16)) : 0xffu);
    if (active && ext_start + n_ext <= (unsigned)L.EX && reg_start + nr <= (unsigned)L.RG)
      sts_u(sb + 4u * ((unsigned)L.oPf + t), (ns ? 1u : 0u) | (np ? 2u : 0u) | (n_ext << 2) | (ext_start << 5) |
                                                 (nr << 10) | (reg_start << 14) | ((unsigned)j << 21));
    n_act += __popc(ab);
    const unsigned tot = __shfl_sync(FULL, incl, 31);
    ext_tot += (int)(tot & 0xffffu);
    reg_tot += (int)(tot >> 16);
  }
  if (__any_sync(FULL, bad != 0) || ext_tot > L.EX || reg_tot > L.RG) return false;
  // ---- chunks of the signals
  int ch_tot = 0;
#pragma unroll 1
  for (int i0 = 0; i0 < L.S2; i0 += 32) {
    const int i = i0 + lane;
    const unsigned c = i < s ? lds_u(sigcnt + 4u * (unsigned)i) : 0u;
    const unsigned nch = (c + 3u) / 4u;
    unsigned incl = nch;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned y = __shfl_up_sync(FULL, incl, d);
      if (lane >= d) incl += y;
    }
    const unsigned start = (unsigned)ch_tot + incl - nch;
    sts_u(sigcnt + 4u * (unsigned)i, start);
    sts_u(sb + 4u * (unsigned)(L.oCol + i), start | (nch << 16));
    ch_tot += (int)__shfl_sync(FULL, incl, 31);
  }
  if (ch_tot > L.CH) return false;
  __syncwarp();
  // ---- per-protein scalars in active order (asynchronous; awaited by the caller)
  {
    const long long prow = prm_row * (long long)p;
    const float *gv = a.adj_vmax ? a.adj_vmax : a.Vmax;
#pragma unroll 1
    for (int t = lane; t < n_act; t += 32) {
      const unsigned j = lds_u(sb + 4u * (unsigned)(L.oPf + t)) >> 21;
      cp_async4(sb + 4u * (unsigned)(L.oKe + t), a.Ke + prow + j);
      cp_async4(sb + 4u * (unsigned)(L.oKmf + t), a.Kmf + prow + j);
      cp_async4(sb + 4u * (unsigned)(L.oKmb + t), a.Kmb + prow + j);
      cp_async4(sb + 4u * (unsigned)(L.oVmax + t), gv + prow + j);
    }
  }
  // ---- pass B: place the entries
  const long long tile = prm_row * (long long)p * s;
  int rare_tot = 0, reg_run = 0;
#pragma unroll 1
  for (int base = 0; base < nq; base += 32) {
    const int k = base + lane;
    const bool v = k < nq;
    const uint2 qw = v ? lds_v2(aQueue + 8u * (unsigned)k) : make_uint2(0u, 0u);
    const unsigned sig = qw.x & 0xffu, row = (qw.x >> 8) & 0xffu;
    const int av = (int)(int8_t)((qw.x >> 16) & 0xffu);
    const unsigned nf = qw.y & 0xffu, nb = (qw.y >> 8) & 0xffu;
    const int n = (int)qw.y >> 16;
    const unsigned rm = v ? lds_u(rowmap + 4u * row) : 0xffu;
    const bool active = (rm & 0xffu) != 0xffu;
    const unsigned t = rm & 0xffu, ext_start = (rm >> 8) & 0xffu;
    // ranks inside the row (the row's elements are consecutive, ascending signal)
    const unsigned peers = __match_any_sync(FULL, v ? row : 0x100u + (unsigned)lane);
    const unsigned bs = __ballot_sync(FULL, nf != 0u), bp = __ballot_sync(FULL, nb != 0u);
    const unsigned run = v ? lds_u(rowrun + 4u * row) : 0u;
    const unsigned ks = (run & 0xffu) + __popc(peers & bs & lt), kp = (run >> 8) + __popc(peers & bp & lt);
    // table slots of exponents >= 4, in queue order (substrate side first)
    const unsigned rs = __ballot_sync(FULL, nf > 3u), rp = __ballot_sync(FULL, nb > 3u);
    const unsigned r_sub = (unsigned)rare_tot + __popc(rs & lt) + __popc(rp & lt);
    const unsigned r_prod = r_sub + (nf > 3u ? 1u : 0u);
    if (nf) {
      unsigned idx = 1u + (nf - 1u) * (unsigned)L.S2 + sig;
      if (nf > 3u) {
        idx = 1u + 3u * (unsigned)L.S2 + r_sub;
        sts_u(sb + 4u * ((unsigned)L.oRare + r_sub), sig | (nf << 8));
      }
      const unsigned word = ks < 4u ? (unsigned)L.oEnt8 + 2u * t : (unsigned)L.oExt8 + 2u * (ext_start + ks / 4u - 1u);
      sts_u8(sb + 4u * word + (ks & 3u), idx);
    }
    if (nb) {
      unsigned idx = 1u + (nb - 1u) * (unsigned)L.S2 + sig;
      if (nb > 3u) {
        idx = 1u + 3u * (unsigned)L.S2 + r_prod;
        sts_u(sb + 4u * ((unsigned)L.oRare + r_prod), sig | (nb << 8));
      }
      const unsigned word = kp < 4u ? (unsigned)L.oEnt8 + 2u * t : (unsigned)L.oExt8 + 2u * (ext_start + kp / 4u - 1u);
      sts_u8(sb + 4u * (word + 1u) + (kp & 3u), idx);
    }
    rare_tot += __popc(rs) + __popc(rp);
    // regulatory entries of active proteins, in queue order = grouped by protein, ascending signal
    const unsigned br = __ballot_sync(FULL, av != 0 && active);
    if (av != 0 && active) {
      const unsigned pos = (unsigned)reg_run + __popc(br & lt);
      sts_u(sb + 4u * ((unsigned)L.oRegw + pos), sig | ((unsigned)(av & 0xff) << 8));
      cp_async4(sb + 4u * ((unsigned)L.oRegk + pos), a.Kmr + tile + (long long)row * s + sig);
    }
    reg_run += __popc(br);
    // N entries: per signal in ascending protein order (queue order).  Entries of inactive proteins
    // (N != 0 without Nf / Nb) have V = 0: they would add +-0 and are left out like in run_sparse,
    // whose csc holds them but multiplies by V = 0.
    const bool has_n = v && n != 0;
    const unsigned speers = __match_any_sync(FULL, has_n ? sig : 0x100u + (unsigned)lane);
    if (has_n) {
      const unsigned pos = 4u * lds_u(sigcnt + 4u * sig) + lds_u(sigrun + 4u * sig) + __popc(speers & lt);
      const unsigned tt = active ? t : (unsigned)L.PA;
      sts_u16r(sb + 4u * (unsigned)L.oCsc + 2u * pos, tt | (((unsigned)n & 0x1ffu) << 7));
    }
    __syncwarp();
    if (v && lane == 31 - __clz(peers)) sts_u(rowrun + 4u * row, run + (__popc(peers & bs) | (__popc(peers & bp) << 8)));
    if (has_n && lane == 31 - __clz(speers)) red_add_shared(sigrun + 4u * sig, (unsigned)__popc(speers));
    __syncwarp();
  }
  cnt.n_act = n_act;
  cnt.n_rare = rare_tot;
  cnt.n_reg = reg_run;
  return true;
}

// powers of one signal into the table (same multiplications as powi / entry_factor)
__device__ __forceinline__ void table_store(unsigned aT, unsigned S2x4, unsigned i, float x) {
  const float x2 = x * x;
  sts_f(aT + 4u + 4u * i, x);
  sts_f(aT + 4u + S2x4 + 4u * i, x2);
  sts_f(aT + 4u + 2u * S2x4 + 4u * i, x * x2);
}

// rare slots: X_sig^n for the entries with n >= 4
__device__ __noinline__ void table_rare(unsigned aT, unsigned S2x4, unsigned aRare, unsigned aX, int n_rare, int lane) {
#pragma unroll 1
  for (int r = lane; r < n_rare; r += 32) {
    const unsigned w = lds_u(aRare + 4u * (unsigned)r);
    sts_f(aT + 4u + 3u * S2x4 + 4u * (unsigned)r, powi(lds_f(aX + 4u * (w & 0xffu)), (int)(w >> 8)));
  }
}

// products over a protein's substrate and product entries (kinetics.py:894-918) from the table
__device__ MS_PRODUCTS_INLINE void lane_products(unsigned aT, unsigned aEnt8, unsigned aExt8, unsigned t, unsigned pf,
                                              float &xf, float &xb) {
  uint2 e = lds_v2(aEnt8 + 8u * t);
  // byte k of a word -> table address: one PRMT and one shift-add
#define MS_T(word, k) lds_f(aT + (__byte_perm(word, 0u, 0x4440u | (k)) << 2))
  xf = MS_T(e.x, 0);
  xb = MS_T(e.y, 0);
  xf = xf * MS_T(e.x, 1);
  xb = xb * MS_T(e.y, 1);
  xf = xf * MS_T(e.x, 2);
  xb = xb * MS_T(e.y, 2);
  xf = xf * MS_T(e.x, 3);
  xb = xb * MS_T(e.y, 3);
  unsigned n_ext = (pf >> 2) & 7u;
  if (n_ext) {  // more than 4 entries on a side (3 % of the proteins)
    unsigned ax = aExt8 + 8u * ((pf >> 5) & 31u);
#pragma unroll 1
    do {
      e = lds_v2(ax);
      xf = xf * MS_T(e.x, 0);
      xb = xb * MS_T(e.y, 0);
      xf = xf * MS_T(e.x, 1);
      xb = xb * MS_T(e.y, 1);
      xf = xf * MS_T(e.x, 2);
      xb = xb * MS_T(e.y, 2);
      xf = xf * MS_T(e.x, 3);
      xb = xb * MS_T(e.y, 3);
      ax += 8u;
    } while (--n_ext);
  }
#undef MS_T
  xf = fix_product(xf);
  xb = fix_product(xb);
}

// V of every protein of a cell, 0 for those without entries (single-part calls; kept out of line)
__device__ __noinline__ void store_velocities(float *vrow, int p, int lane, unsigned aPf, unsigned aV, int n_act) {
#pragma unroll 1
  for (int j = lane; j < p; j += 32) vrow[j] = 0.0f;
  __syncwarp();
#pragma unroll 1
  for (int t = lane; t < n_act; t += 32) vrow[lds_u(aPf + 4u * (unsigned)t) >> 21] = lds_f(aV + 4u * (unsigned)t);
}

__device__ __forceinline__ int csc_n(unsigned h) { return (int)(short)(unsigned short)h >> 7; }

// ---------------------------------------------------------------------------
// all parts of one cell on the fixed-format lists; one warp, no CTA barrier
// ---------------------------------------------------------------------------
__device__ void run_lanes(const KinArgs &a, const LaneLayout &L, unsigned sb, long long cell, int lane,
                          const LaneCounts cnt, unsigned &wflags) {
  const int s = a.s;
  const float qnan = __int_as_float(0x7fc00000);
  const unsigned aCol = sb + 4u * L.oCol, aKe = sb + 4u * L.oKe, aKmf = sb + 4u * L.oKmf, aKmb = sb + 4u * L.oKmb;
  const unsigned aVmax = sb + 4u * L.oVmax, aPf = sb + 4u * L.oPf, aEnt8 = sb + 4u * L.oEnt8, aExt8 = sb + 4u * L.oExt8;
  const unsigned aRare = sb + 4u * L.oRare, aRegw = sb + 4u * L.oRegw, aRegk = sb + 4u * L.oRegk, aCsc = sb + 4u * L.oCsc;
  const unsigned aX0 = sb + 4u * L.oX0, aX1 = sb + 4u * L.oX1, aRegf = sb + 4u * L.oRegf, aNv = sb + 4u * L.oNv;
  const unsigned aT = sb + 4u * L.oT, aV = sb + 4u * L.oV, aF = sb + 4u * L.oF, aFk = sb + 4u * L.oFk;
  const unsigned S2x4 = 4u * (unsigned)L.S2;
  const int n_act = cnt.n_act, n_rare = cnt.n_rare, n_reg = cnt.n_reg;
  if (lane == 0) sts_f(aT, 1.0f);  // the neutral entry
#pragma unroll 1
  for (int q = 0; q < a.n_parts; q++) {
    const float trim = a.trims[q];
    // table of X0, regulatory factors (kinetics.py:789-799), entry state of the part
    if (q == 0) {  // later parts start from the previous part's X1, whose powers the table already holds
#pragma unroll 1
      for (int i = lane; i < s; i += 32) table_store(aT, S2x4, (unsigned)i, lds_f(aX0 + 4u * (unsigned)i));
      if (n_rare) table_rare(aT, S2x4, aRare, aX0, n_rare, lane);
    }
#pragma unroll 1
    for (int e = lane; e < n_reg; e += 32) {
      const unsigned wd = lds_u(aRegw + 4u * (unsigned)e);
      sts_f(aRegf + 4u * (unsigned)e, reg_factor(lds_f(aX0 + 4u * (wd & 0xffu)), (int)(int8_t)((wd >> 8) & 0xff),
                                                 lds_f(aRegk + 4u * (unsigned)e)));
    }
#pragma unroll 1
    for (int t = lane; t < L.PA2; t += 32) {
      sts_f(aV + 4u * (unsigned)t, 0.0f);  // incl. the dummy protein of the padding entries
      sts_f(aF + 4u * (unsigned)t, 1.0f);
      sts_u(aFk + 4u * (unsigned)t, KEY_ONE);
    }
    __syncwarp();
    // ---- velocities (kinetics.py:771-806)
    bool poison = false;
#pragma unroll 1
    for (int t = lane; t < n_act; t += 32) {
      const unsigned pf = lds_u(aPf + 4u * (unsigned)t);
      float xf, xb;
      lane_products(aT, aEnt8, aExt8, (unsigned)t, pf, xf, xb);
      float kf = xf / lds_f(aKmf + 4u * (unsigned)t);
      if (!(pf & 1u)) kf = 0.0f;
      if (isinf(kf)) kf = V_MAX;
      float kb = xb / lds_f(aKmb + 4u * (unsigned)t);
      if (!(pf & 2u)) kb = 0.0f;
      if (isinf(kb)) kb = V_MAX;
      float a_reg = 1.0f;
      unsigned nr = (pf >> 10) & 15u;
      if (nr) {
        unsigned ar = aRegf + 4u * ((pf >> 14) & 127u);
#pragma unroll 1
        do {
          a_reg = a_reg * lds_f(ar);
          ar += 4u;
        } while (--nr);
      }
      float vmax = lds_f(aVmax + 4u * (unsigned)t);
      if (!a.adj_vmax) vmax = clamp0(vmax * trim);  // kinetics.py:749
      const float v = velocity_from(kf, kb, a_reg, vmax);
      sts_f(aV + 4u * (unsigned)t, v);
      poison |= isnan(v);
    }
    poison = __any_sync(FULL, poison);  // 0 * NaN of the dense reference reaches every signal
    __syncwarp();
    // ---- substrate trimming (kinetics.py:861-879)
#pragma unroll 1
    for (int i = lane; i < s; i += 32) {
      const unsigned cw = lds_u(aCol + 4u * (unsigned)i);
      const unsigned c0 = cw & 0xffffu, nc = cw >> 16;
      float d = 0.0f;
#pragma unroll 1
      for (unsigned c = c0; c < c0 + nc; c++) {
        const uint2 w = lds_v2(aCsc + 8u * c);
        float4 nv;
        nv.x = (float)csc_n(w.x) * lds_f(aV + ((w.x & 0x7fu) << 2));          // kinetics.py:759
        nv.y = (float)csc_n(w.x >> 16) * lds_f(aV + ((w.x >> 14) & 0x1fcu));
        nv.z = (float)csc_n(w.y) * lds_f(aV + ((w.y & 0x7fu) << 2));
        nv.w = (float)csc_n(w.y >> 16) * lds_f(aV + ((w.y >> 14) & 0x1fcu));
        sts_f4(aNv + 16u * c, nv);
        float neg;
        neg = -nv.x; d = d + (neg < 0.0f ? 0.0f : neg);
        neg = -nv.y; d = d + (neg < 0.0f ? 0.0f : neg);
        neg = -nv.z; d = d + (neg < 0.0f ? 0.0f : neg);
        neg = -nv.w; d = d + (neg < 0.0f ? 0.0f : neg);
      }
      float f = lds_f(aX0 + 4u * (unsigned)i) / d;
      if (f > 1.0f) f = 1.0f;
      const unsigned key = fkey(f);
#pragma unroll 1
      for (unsigned c = c0; c < c0 + nc; c++) {  // every protein that removes signal i is limited by F_i
        const uint2 w = lds_v2(aCsc + 8u * c);
        const float4 nv = lds_f4(aNv + 16u * c);
        if (nv.x < 0.0f) atom_min_shared(aFk + ((w.x & 0x7fu) << 2), key);
        if (nv.y < 0.0f) atom_min_shared(aFk + ((w.x >> 14) & 0x1fcu), key);
        if (nv.z < 0.0f) atom_min_shared(aFk + ((w.y & 0x7fu) << 2), key);
        if (nv.w < 0.0f) atom_min_shared(aFk + ((w.y >> 14) & 0x1fcu), key);
      }
    }
    __syncwarp();
    // the limiting factor of each protein as a float (decoded once per protein, not once per entry)
#pragma unroll 1
    for (int t = lane; t < L.PA2; t += 32) sts_f(aFk + 4u * (unsigned)t, fkey_decode(lds_u(aFk + 4u * (unsigned)t)));
    __syncwarp();
    // X1 = max(X0 + sum_j NV'_ji, 0)   (kinetics.py:763-764)
#pragma unroll 1
    for (int i = lane; i < s; i += 32) {
      const unsigned cw = lds_u(aCol + 4u * (unsigned)i);
      const unsigned c0 = cw & 0xffffu, nc = cw >> 16;
      float acc = 0.0f;
#pragma unroll 1
      for (unsigned c = c0; c < c0 + nc; c++) {
        const uint2 w = lds_v2(aCsc + 8u * c);
        float4 nv = lds_f4(aNv + 16u * c);
        nv.x = nv.x * lds_f(aFk + ((w.x & 0x7fu) << 2));
        nv.y = nv.y * lds_f(aFk + ((w.x >> 14) & 0x1fcu));
        nv.z = nv.z * lds_f(aFk + ((w.y & 0x7fu) << 2));
        nv.w = nv.w * lds_f(aFk + ((w.y >> 14) & 0x1fcu));
        sts_f4(aNv + 16u * c, nv);
        acc = acc + nv.x;
        acc = acc + nv.y;
        acc = acc + nv.z;
        acc = acc + nv.w;
      }
      float x = clamp0(lds_f(aX0 + 4u * (unsigned)i) + acc);
      if (poison) x = qnan;
      sts_f(aX1 + 4u * (unsigned)i, x);
      table_store(aT, S2x4, (unsigned)i, x);
    }
    __syncwarp();
    if (n_rare) {
      table_rare(aT, S2x4, aRare, aX1, n_rare, lane);
      __syncwarp();
    }

    // ---- equilibrium trimming (kinetics.py:808-859)
    const int iters = a.eq_iters[q];
    float inc = 1.0f;
#pragma unroll 1
    for (int k = 0; k < iters; k++) {
      if (k < 4) inc = inc * 0.5f;  // 0.5, 0.25, 0.125, 0.0625 (kinetics.py:846)
      unsigned mine = 0;
      bool changed = false;
#pragma unroll 1
      for (int t = lane; t < n_act; t += 32) {
        const unsigned pf = lds_u(aPf + 4u * (unsigned)t);
        float xf, xb;
        lane_products(aT, aEnt8, aExt8, (unsigned)t, pf, xf, xb);
        if (!(pf & 1u)) xf = 0.0f;
        if (!(pf & 2u)) xb = 0.0f;
        const float F = lds_f(aF + 4u * (unsigned)t);
        const unsigned dec = eq_decide_fast(lds_f(aV + 4u * (unsigned)t), quotient_fixed(xb, xf),
                                            lds_f(aKe + 4u * (unsigned)t), F);
        if (dec & 3u) {
          mine |= (dec & 4u) ? 3u : 2u;
          const float Fn = eq_update(F, dec, inc);
          changed |= (Fn != F);
          sts_f(aF + 4u * (unsigned)t, Fn);
        }
      }
      const unsigned any = __reduce_or_sync(FULL, mine);
      // If no F moved, every later iteration of this part repeats this one exactly (run_sparse).
      const bool moved = __any_sync(FULL, changed);
      if (any) {  // flag bits q*4+kk and 12+q*4+kk for kk in [k, k_end), warp-uniform
        const int k_end = moved ? k + 1 : iters;
        const unsigned span = ((1u << k_end) - (1u << k)) << (4 * q);
        wflags |= (span << 12) | ((any & 1u) ? span : 0u);
      }
      if (!moved) break;
      __syncwarp();
#pragma unroll 1
      for (int i = lane; i < s; i += 32) {
        const unsigned cw = lds_u(aCol + 4u * (unsigned)i);
        const unsigned c0 = cw & 0xffffu, nc = cw >> 16;
        float acc = 0.0f;
#pragma unroll 1
        for (unsigned c = c0; c < c0 + nc; c++) {
          const uint2 w = lds_v2(aCsc + 8u * c);
          const float4 nv = lds_f4(aNv + 16u * c);
          acc = acc + nv.x * lds_f(aF + ((w.x & 0x7fu) << 2));
          acc = acc + nv.y * lds_f(aF + ((w.x >> 14) & 0x1fcu));
          acc = acc + nv.z * lds_f(aF + ((w.y & 0x7fu) << 2));
          acc = acc + nv.w * lds_f(aF + ((w.y >> 14) & 0x1fcu));
        }
        float x = clamp0(lds_f(aX0 + 4u * (unsigned)i) + acc);
        if (poison) x = qnan;
        sts_f(aX1 + 4u * (unsigned)i, x);
        table_store(aT, S2x4, (unsigned)i, x);
      }
      __syncwarp();
      if (n_rare) {
        table_rare(aT, S2x4, aRare, aX1, n_rare, lane);
        __syncwarp();
      }
    }

    if (a.V_out) store_velocities(a.V_out + cell * (long long)a.p, a.p, lane, aPf, aV, n_act);  // single-part calls only
    // next part starts from this part's result
#pragma unroll 1
    for (int i = lane; i < s; i += 32) sts_f(aX0 + 4u * (unsigned)i, lds_f(aX1 + 4u * (unsigned)i));
    __syncwarp();
  }
}

}  // namespace ms
