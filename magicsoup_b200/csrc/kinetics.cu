// Fused per-cell kinetics for sm_100a: Kinetics.integrate_signals
// (reference python/magicsoup/kinetics.py:725-918) and the gather half of
// World.enzymatic_activity (python/magicsoup/world.py:619-622).
//
// One warp integrates one cell.  The cell's dense (p,s) parameter tile is streamed
// from HBM exactly once with coalesced loads (one 128 B row segment per load
// instruction) and compressed on the fly into sparse lists in shared memory:
//   cat  : per protein, ascending signal: (signal, Nf, Nb, N)      -> products X^N
//   reg  : per protein, ascending signal: (signal, A) + Kmr        -> allosteric term
//   csc  : per signal, ascending protein: (protein, N)             -> sums over proteins
// All 3 parts x (velocities, substrate trimming, <=4 equilibrium iterations) then run
// on those lists out of shared memory.  Cells whose tiles do not fit the lists or the
// 8-bit packing take a dense path straight from global memory with identical
// arithmetic.  See DESIGN.md for the layout, the spec order and the roofline.
#include <stdlib.h>

#include "ms_common.cuh"
#include "kinetics_layout.cuh"

namespace ms {

constexpr int WARPS_PER_CTA = 8;
constexpr int MAX_THREADS = 32 * WARPS_PER_CTA;  // every launch uses at most 8 warps per CTA
#ifndef MS_TUNE_MIN_CTAS
#define MS_TUNE_MIN_CTAS 4
#endif
constexpr int MAX_S_SPARSE = 256;  // 8-bit signal / protein indices in the packed lists
constexpr unsigned FULL = 0xffffffffu;

struct KinArgs {
  const int *N, *Nf, *Nb, *A;
  const float *Kmr, *Ke, *Kmf, *Kmb, *Vmax;
  long long c;
  int p, s;
  // signal source: X_in (c,s), or gather from cell_molecules / molecule_map
  const float *X_in;
  const float *cell_mols;
  const float *mol_map;
  const int *pos;
  int m, H, W;
  float *X_out;
  unsigned *flags;
  int n_parts;            // 3 = integrate_signals, 1 = one part with adj_vmax
  int eq_iters[3];
  float trims[3];
  const float *adj_vmax;  // (c,p) when n_parts == 1
  float *V_out;           // optional (c,p), n_parts == 1 only
  int e_ent, e_nnz, e_reg;  // list capacities (entries)
  int words_per_warp;     // shared memory words per warp
  unsigned region_bytes;  // the same in bytes
  int smem_align;         // alignment of the warp regions in bytes (power of two)
  int sparse_enabled;     // 0 forces the dense path (p > 256 or s > 128)
  // Two tiers.  Tier 1 (small lists, 32 warps/SM) defers cells whose tile does not fit to
  // defer_list; tier 2 (lists sized for a full tile, few warps/SM) integrates exactly those
  // cells, and only what cannot be packed at all takes the dense path.
  int tier;               // 1 or 2
  int defer_enabled;      // tier 1: a tier-2 launch follows
  int *defer_count;       // device counter (zeroed before tier 1)
  int *defer_list;        // device list of deferred cell indices (capacity c)
  int *work_counter;      // device counter handing out cells to warps (zeroed before the launch)
  int scan_staged;        // batch mode: steps of the tile scan staged ahead in shared memory (0, 1 or 2)
  int batch_enabled;      // tier 1, p, s <= 64, 8 warps per CTA: the CTA integrates 8 cells together
  LaneLayout lanes;       // region layout of the lane kernel (host-computed: constant-bank operands)
  // Parameter rows and the list cache (ms_*_cached entry points).  row_map[cell] is the row of the
  // nine parameter tensors (and of the cache) that holds logical cell `cell`; X, positions and cell
  // molecules are always indexed by the logical cell.  cache: one record per parameter row with the
  // lists build_lists produced for it (batch mode only); rec_seg: the three region segments a record
  // holds, {word offset, words} each.
  const int *row_map;
  unsigned *phase_recs;   // phased lane kernel: (grid x phase_batch) records of scratch, see kinetics_lanes_phased_kernel
  int phase_batch, phase_ahead;
  unsigned *cache;
  int rec_words;
  int rec_seg[6];
};

// list-cache record: word 0 = n_act | status << 16, words 1..3 unused, then the three segments
constexpr unsigned REC_STALE = 0, REC_VALID = 1, REC_OVERFLOW = 2;  // in bits 24.. of word 0
constexpr int REC_HEADER_WORDS = 4;

__device__ __forceinline__ long long param_row(const KinArgs &a, long long cell) {
  return a.row_map ? (long long)a.row_map[cell] : cell;
}

// ---------------------------------------------------------------------------
// Per-warp shared memory region (32-bit words).  The arrays that the hot loops index with
// a value taken from a packed list entry (signal -> X0/X1, protein -> V, F, ...) start at
// addresses aligned to their power-of-two stride, so that `base | (entry & 0x3fc)` is the
// element's address (one LOP3, no shift/add).  The hot loops address shared memory with
// explicit 32-bit shared-window addresses.
//
//   X0 [S2] X1 [S2] Fsig/cursor [S2] colptr [S2+32] pad
//   V F FminKey Ke Kmf Kmb Vmax act [P2 each]  rowptr [P2+32]
//   ent [e_ent]  nvs [e_nnz] csc [e_nnz]  reg [e_reg] regk [e_reg]
//
// ent : signal*4 | exponent << 10 | (1 << 31 if product side); per protein, ascending signal
// csc : protein*4 | N << 10; per signal, ascending protein; nvs: the rate of that entry
// act : protein*4 | (has substrates) << 10 | (has products) << 11
// rowptr : ent offset | reg offset << 16;  reg: signal | (A & 0xff) << 8 | protein << 16
// During the scan nvs+csc hold the queue of non-zero tile elements (2 words each) and
// then (nvs) the unsorted N list, which is sorted into csc.
// ---------------------------------------------------------------------------

struct Layout {
  int S2, P2, oP, oEnt, oCsc, oNvs, oReg, oRegk, words;
};

__host__ __device__ inline Layout make_layout(int p, int s, int e_ent, int e_nnz, int e_reg) {
  Layout L;
  L.S2 = pow2_at_least_32(s);
  L.P2 = pow2_at_least_32(p);
  L.oP = ((4 * L.S2 + 32 + L.P2 - 1) / L.P2) * L.P2;
  L.oEnt = L.oP + 9 * L.P2 + 32;
  L.oNvs = L.oEnt + e_ent;
  L.oCsc = L.oNvs + e_nnz;
  L.oReg = L.oCsc + e_nnz;
  L.oRegk = L.oReg + e_reg;
  const int align = L.S2 > L.P2 ? L.S2 : L.P2;
  L.words = ((L.oRegk + e_reg + align - 1) / align) * align;
  return L;
}

struct WarpMem {
  unsigned *b;    // generic pointer to the region
  unsigned sb;    // the same location as a 32-bit shared-window byte address
  int S2, P2, oP, oEnt, oCsc, oNvs, oReg, oRegk;
  // generic views (scan / list building)
  __device__ __forceinline__ float *X0() const { return (float *)b; }
  __device__ __forceinline__ unsigned *cursor() const { return b + 2 * S2; }
  __device__ __forceinline__ unsigned *colptr() const { return b + 3 * S2; }
  __device__ __forceinline__ unsigned *FminKey() const { return b + oP + 2 * P2; }
  __device__ __forceinline__ float *Ke() const { return (float *)(b + oP + 3 * P2); }
  __device__ __forceinline__ float *Kmf() const { return (float *)(b + oP + 4 * P2); }
  __device__ __forceinline__ float *Kmb() const { return (float *)(b + oP + 5 * P2); }
  __device__ __forceinline__ float *Vmax() const { return (float *)(b + oP + 6 * P2); }
  __device__ __forceinline__ unsigned *act() const { return b + oP + 7 * P2; }
  __device__ __forceinline__ unsigned *rowptr() const { return b + oP + 8 * P2; }
  __device__ __forceinline__ unsigned *ent() const { return b + oEnt; }
  __device__ __forceinline__ unsigned *csc() const { return b + oCsc; }
  __device__ __forceinline__ unsigned *nvs() const { return b + oNvs; }
  __device__ __forceinline__ unsigned *queue() const { return b + oNvs; }  // nvs + csc: 2 * e_nnz words
  __device__ __forceinline__ unsigned *reg() const { return b + oReg; }
  __device__ __forceinline__ float *regk() const { return (float *)(b + oRegk); }
  // shared-window byte addresses (hot loops)
  __device__ __forceinline__ unsigned aX0() const { return sb; }
  __device__ __forceinline__ unsigned aX1() const { return sb + 4 * S2; }
  __device__ __forceinline__ unsigned aCol() const { return sb + 12 * S2; }
  __device__ __forceinline__ unsigned aV() const { return sb + 4 * oP; }
  __device__ __forceinline__ unsigned aF() const { return sb + 4 * (oP + P2); }
  __device__ __forceinline__ unsigned aFk() const { return sb + 4 * (oP + 2 * P2); }
  __device__ __forceinline__ unsigned aKe() const { return sb + 4 * (oP + 3 * P2); }
  __device__ __forceinline__ unsigned aKmf() const { return sb + 4 * (oP + 4 * P2); }
  __device__ __forceinline__ unsigned aKmb() const { return sb + 4 * (oP + 5 * P2); }
  __device__ __forceinline__ unsigned aVmax() const { return sb + 4 * (oP + 6 * P2); }
  __device__ __forceinline__ unsigned aAct() const { return sb + 4 * (oP + 7 * P2); }
  __device__ __forceinline__ unsigned aRow() const { return sb + 4 * (oP + 8 * P2); }
  __device__ __forceinline__ unsigned aEnt() const { return sb + 4 * oEnt; }
  __device__ __forceinline__ unsigned aCsc() const { return sb + 4 * oCsc; }
  __device__ __forceinline__ unsigned aNvs() const { return sb + 4 * oNvs; }
  __device__ __forceinline__ unsigned aReg() const { return sb + 4 * oReg; }
  __device__ __forceinline__ unsigned aRegk() const { return sb + 4 * oRegk; }
};

__device__ __forceinline__ WarpMem carve(unsigned *w, unsigned shared_addr, int p, int s, int e_ent,
                                         int e_nnz, int e_reg) {
  const Layout L = make_layout(p, s, e_ent, e_nnz, e_reg);
  WarpMem m;
  m.b = w; m.sb = shared_addr;
  m.S2 = L.S2; m.P2 = L.P2; m.oP = L.oP; m.oEnt = L.oEnt; m.oCsc = L.oCsc; m.oNvs = L.oNvs;
  m.oReg = L.oReg; m.oRegk = L.oRegk;
  return m;
}

// explicit shared-memory accesses by 32-bit shared-window address
__device__ __forceinline__ unsigned lds_u(unsigned addr) {
  unsigned v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds_f(unsigned addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u(unsigned addr, unsigned v) {
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_f(unsigned addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ unsigned lds_u16(unsigned addr) {
  unsigned v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void red_or_shared(unsigned addr, unsigned v) {
  asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void atom_min_shared(unsigned addr, unsigned v) {
  asm volatile("red.shared.min.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// asynchronous 4-byte global -> shared copies (LDGSTS): no registers, latency overlaps the scan
__device__ __forceinline__ void cp_async4(unsigned saddr, const void *gptr) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(saddr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// order-preserving unsigned key of a float; NaN maps to 0 so that it wins the atomic min
// (torch.min propagates NaN, kinetics.py:877)
__device__ __forceinline__ unsigned fkey(float f) {
  if (isnan(f)) return 0u;
  const unsigned b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float fkey_decode(unsigned k) {
  if (k == 0u) return __int_as_float(0x7fc00000);
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
constexpr unsigned KEY_ONE = 0xbf800000u;  // fkey(1.0f)

template <int VEC>
struct TileVec {  // VEC consecutive elements of the four integer tiles
  int nf[VEC], nb[VEC], nn[VEC], na[VEC];
};

template <int VEC>
__device__ __forceinline__ void load_vec(const int *ptr, bool ok, int (&v)[VEC]) {
  if (!ok) {
#pragma unroll
    for (int e = 0; e < VEC; e++) v[e] = 0;
    return;
  }
  if constexpr (VEC == 4) {
    const int4 t = __ldcs(reinterpret_cast<const int4 *>(ptr));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  } else if constexpr (VEC == 2) {
    const int2 t = __ldcs(reinterpret_cast<const int2 *>(ptr));
    v[0] = t.x; v[1] = t.y;
  } else {
    v[0] = __ldcs(ptr);
  }
}

template <int VEC>
__device__ __forceinline__ int pick(const int (&v)[VEC], int e) {
  if constexpr (VEC == 4) {
    const int lo = (e & 1) ? v[1] : v[0], hi = (e & 1) ? v[3] : v[2];
    return (e & 2) ? hi : lo;
  } else if constexpr (VEC == 2) {
    return (e & 1) ? v[1] : v[0];
  } else {
    return v[0];
  }
}

// bit e: element e of the lane's VEC elements has a non-zero Nf, Nb, N or A
template <int VEC>
__device__ __forceinline__ unsigned elem_mask(const TileVec<VEC> &t) {
  unsigned mask = 0;
#pragma unroll
  for (int e = 0; e < VEC; e++) mask |= ((t.nf[e] | t.nb[e] | t.nn[e] | t.na[e]) != 0 ? 1u : 0u) << e;
  return mask;
}

// Pushes the non-zero elements of one 32*VEC-element step (lane-major, i.e. row-major order).
__device__ __forceinline__ void queue_store(unsigned *queue, int pos, unsigned w0, unsigned w1) {
  *reinterpret_cast<uint2 *>(queue + 2 * pos) = make_uint2(w0, w1);
}
__device__ __forceinline__ void queue_store(unsigned queue_saddr, int pos, unsigned w0, unsigned w1) {
  asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(queue_saddr + 8u * (unsigned)pos), "r"(w0), "r"(w1) : "memory");
}

template <int VEC, typename Q>
__device__ __forceinline__ void push_step(const TileVec<VEC> &t, unsigned mask, Q queue, int cap, int row,
                                          int sig0, unsigned lt, int &nq, int &chk) {
  if (__ballot_sync(FULL, mask != 0) == 0) return;
  int pos = nq;
#pragma unroll
  for (int e = 0; e < VEC; e++) {
    const unsigned be = __ballot_sync(FULL, (mask >> e) & 1u);
    pos += __popc(be & lt);
    nq += __popc(be);
  }
#pragma unroll 1
  while (mask) {  // most lanes hold no element, the others one or (rarely) more
    const int e = __ffs(mask) - 1;
    mask &= mask - 1u;
    const int nf = pick<VEC>(t.nf, e), nb = pick<VEC>(t.nb, e), nn = pick<VEC>(t.nn, e), na = pick<VEC>(t.na, e);
    // packing: 0 <= Nf, Nb < 256, |N| < 2^15, |A| < 2^7
    chk |= ((nf | nb) & 0xffffff00) | ((nn + 32768) & 0xffff0000) | ((na + 128) & 0xffffff00);
    if (pos < cap)
      queue_store(queue, pos, (unsigned)(sig0 + e) | ((unsigned)row << 8) | ((unsigned)(na & 0xff) << 16),
                  ((unsigned)nf & 0xffu) | (((unsigned)nb & 0xffu) << 8) | ((unsigned)nn << 16));
    pos++;
  }
}

// ---------------------------------------------------------------------------
// scan, phase 1: stream the cell's dense tile once (32*VEC consecutive elements per load
// instruction, row-major) and push every element with a non-zero Nf, Nb, N or A onto a
// queue in shared memory (2 words per element, 8-byte aligned), in row-major order.  The
// loads of a step are issued before the previous step is pushed (two register sets).
// Returns the queue length, or -1 if the tile does not fit the queue / the packing.
// ---------------------------------------------------------------------------
template <int VEC>
__device__ __noinline__ int scan_tile(const KinArgs &a, unsigned *queue, int cap, long long tile, int lane) {
  const int s = a.s, total = a.p * s;
  const int *gN = a.N + tile, *gNf = a.Nf + tile, *gNb = a.Nb + tile, *gA = a.A + tile;
  const unsigned lt = (1u << lane) - 1u;
  constexpr int STEP = 32 * VEC;
  const int drow = STEP / s, dsig = STEP - drow * s;
  int q0 = lane * VEC;
  int row = q0 / s, sig0 = q0 - row * s;
  int nq = 0, chk = 0;
  auto fetch = [&](TileVec<VEC> &t, int q) {
    const bool ok = q < total;
    load_vec<VEC>(gNf + q, ok, t.nf);
    load_vec<VEC>(gNb + q, ok, t.nb);
    load_vec<VEC>(gN + q, ok, t.nn);
    load_vec<VEC>(gA + q, ok, t.na);
  };
  auto advance = [&]() {
    q0 += STEP;
    row += drow;
    sig0 += dsig;
    if (sig0 >= s) { sig0 -= s; row++; }
  };
#pragma unroll 1
  for (int base = 0; base < total; base += STEP) {
    TileVec<VEC> t0;
    fetch(t0, q0);
    push_step<VEC, unsigned *>(t0, elem_mask<VEC>(t0), queue, cap, row, sig0, lt, nq, chk);
    advance();
  }
  if (__any_sync(FULL, chk != 0) || nq > cap) return -1;
  return nq;
}

// The same scan with the tile staged through shared memory (VEC = 4): the 4 x 512 bytes of the
// next one or two steps are in flight (cp.async, no registers) while a step is pushed.  The
// staging slots are arrays of the warp's region that are not in use before build_lists:
//   set 0: ent (2 slots), reg + regk, X1 + cursor + colptr
//   set 1: V + F + FminKey, act + rowptr, Ke + Kmf, Kmb + Vmax        (depth 2 only, p > 32)
// (launch_one checks the sizes).  Each lane reads back the 16 bytes it copied itself, so no warp
// synchronisation is needed.
__device__ __forceinline__ void cp_async16(unsigned saddr, const void *gptr, unsigned dep) {
  // dep: a value the copy must wait for (keeps the refill behind the reads of the slot)
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(gptr), "r"(dep) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void lds_vec4(unsigned addr, int (&v)[4]) {
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(addr));
}

template <int DEPTH>
__device__ __noinline__ int scan_tile_staged(const KinArgs &a, unsigned aEnt, unsigned aReg, unsigned aX1,
                                             unsigned aV, unsigned aAct, unsigned aKe, unsigned aKmb,
                                             unsigned queue, int cap, long long tile, int lane) {
  constexpr int depth = DEPTH;
  constexpr int VEC = 4, STEP = 32 * VEC;
  const int s = a.s, total = a.p * s;
  const int *gN = a.N + tile, *gNf = a.Nf + tile, *gNb = a.Nb + tile, *gA = a.A + tile;
  const unsigned lt = (1u << lane) - 1u, lo = 16u * (unsigned)lane;
  // slots of the step being read (c*) and of the other set (o*); equal when depth is 1
  unsigned cNf = aEnt + lo, cNb = cNf + 512u, cN = aReg + lo, cA = aX1 + lo;
  unsigned oNf = cNf, oNb = cNb, oN = cN, oA = cA;
  if (depth == 2) {
    oNf = aV + lo;
    oNb = aAct + lo;
    oN = aKe + lo;
    oA = aKmb + lo;
  }
  const int drow = STEP / s, dsig = STEP - drow * s;
  int q0 = lane * VEC;
  int row = q0 / s, sig0 = q0 - row * s;
  int nq = 0, chk = 0;
  auto issue = [&](int q, unsigned sNf, unsigned sNb, unsigned sN, unsigned sA, unsigned dep) {
    if (q < total) {  // one commit group per step, empty past the end
      cp_async16(sNf, gNf + q, dep);
      cp_async16(sNb, gNb + q, dep);
      cp_async16(sN, gN + q, dep);
      cp_async16(sA, gA + q, dep);
    }
    cp_async_commit();
  };
  issue(q0, cNf, cNb, cN, cA, 0u);
  if (depth == 2) issue(q0 + STEP, oNf, oNb, oN, oA, 0u);
#pragma unroll 1
  for (int base = 0; base < total; base += STEP) {
    if (depth == 2) asm volatile("cp.async.wait_group 1;" ::: "memory");
    else asm volatile("cp.async.wait_group 0;" ::: "memory");
    TileVec<VEC> t;  // past the end of the tile the slots hold stale words: masked out
    lds_vec4(cNf, t.nf);
    lds_vec4(cNb, t.nb);
    lds_vec4(cN, t.nn);
    lds_vec4(cA, t.na);
    const unsigned mask = q0 < total ? elem_mask<VEC>(t) : 0u;
    issue(q0 + depth * STEP, cNf, cNb, cN, cA, mask);  // refill the slots just read
    push_step<VEC, unsigned>(t, mask, queue, cap, row, sig0, lt, nq, chk);
    unsigned x;
    x = cNf; cNf = oNf; oNf = x;
    x = cNb; cNb = oNb; oNb = x;
    x = cN; cN = oN; oN = x;
    x = cA; cA = oA; oA = x;
    q0 += STEP;
    row += drow;
    sig0 += dsig;
    if (sig0 >= s) { sig0 -= s; row++; }
  }
  if (__any_sync(FULL, chk != 0) || nq > cap) return -1;
  return nq;
}

// ---------------------------------------------------------------------------
// scan, phase 2: one lane per queued element: distribute it to the per-protein lists
// (ent, reg) and the unsorted N list, then sort the N list by signal into csc (stable
// counting sort via warp match, so each signal's entries stay in ascending protein
// order).  Returns false if a list overflows.
// ---------------------------------------------------------------------------
template <bool WAIT_ASYNC>
__device__ bool build_lists(const KinArgs &a, const WarpMem &w, long long prm_row, int lane, int nq,
                            int &n_act_out) {
  const int p = a.p, s = a.s;
  const unsigned lt = (1u << lane) - 1u;
  unsigned *queue = w.queue();
  unsigned *tmp = w.queue();       // unsorted N list, written over consumed queue words
  unsigned *rowcnt = w.FminKey();  // per protein: substrate entries | product entries << 10 | reg entries << 20
  int ent_tot = 0, nnz_tot = 0, reg_tot = 0;
  unsigned *cursor = w.cursor();   // per signal: N entries, later the write position of the sort
  for (int j = lane; j < p; j += 32) rowcnt[j] = 0;
  for (int i = lane; i < s; i += 32) cursor[i] = 0;
  __syncwarp();
#pragma unroll 1
  for (int base = 0; base < nq; base += 32) {
    const int k = base + lane;
    const bool v = k < nq;
    const uint2 qw = v ? *reinterpret_cast<const uint2 *>(queue + 2 * k) : make_uint2(0u, 0u);
    const unsigned w0 = qw.x, w1 = qw.y;
    __syncwarp();  // the whole chunk is in registers before tmp overwrites queue words
    const unsigned sig = w0 & 0xffu, row = (w0 >> 8) & 0xffu;
    const int av = (int)(int8_t)((w0 >> 16) & 0xffu);
    const unsigned nf = w1 & 0xffu, nb = (w1 >> 8) & 0xffu;
    const int n = (int)w1 >> 16;
    const unsigned bs = __ballot_sync(FULL, nf != 0), bp = __ballot_sync(FULL, nb != 0);
    const unsigned bn = __ballot_sync(FULL, n != 0), br = __ballot_sync(FULL, av != 0);
    // entries of one element: substrate side first, then product side
    int pe = ent_tot + __popc(bs & lt) + __popc(bp & lt);
    const int pr = reg_tot + __popc(br & lt);
    const int pn = nnz_tot + __popc(bn & lt);
    const unsigned inc = (nf ? 1u : 0u) | (nb ? 1u << 10 : 0u) | (av ? 1u << 20 : 0u);
    if (inc) atomicAdd(&rowcnt[row], inc);
    if (nf) {
      if (pe < a.e_ent) w.ent()[pe] = (sig << 2) | (nf << 10);
      pe++;
    }
    if (nb && pe < a.e_ent) w.ent()[pe] = (sig << 2) | (nb << 10) | 0x80000000u;
    if (n != 0) {
      atomicAdd(&cursor[sig], 1u);
      if (pn < a.e_nnz) tmp[pn] = sig | (row << 8) | ((unsigned)n << 16);
    }
    if (av != 0 && pr < a.e_reg) w.reg()[pr] = sig | ((unsigned)(av & 0xff) << 8) | (row << 16);
    ent_tot += __popc(bs) + __popc(bp);
    nnz_tot += __popc(bn);
    reg_tot += __popc(br);
  }
  if (ent_tot > a.e_ent || nnz_tot > a.e_nnz || reg_tot > a.e_reg) return false;
  __syncwarp();
  {  // Kmr of the (few) regulatory entries: asynchronous gather, overlapped with the sort below
    const long long tile = prm_row * (long long)p * s;
    for (int e = lane; e < reg_tot; e += 32) {
      const unsigned wd = w.reg()[e];
      cp_async4(w.aRegk() + 4 * e, a.Kmr + tile + (long long)(wd >> 16) * s + (wd & 0xffu));
    }
  }

  // row pointers (the lists are in row order, so a row starts where the counts of the rows before
  // it end) and the proteins with substrate/product entries, in ascending order
  int n_act = 0;
  unsigned row_carry = 0;  // ent entries | reg entries << 16 before this chunk of rows
#pragma unroll 1
  for (int j0 = 0; j0 < p; j0 += 32) {
    const int j = j0 + lane;
    const unsigned cw = j < p ? rowcnt[j] : 0u;
    const unsigned cf = cw & 0x3ffu, cb = (cw >> 10) & 0x3ffu;
    const unsigned cnt = (cf + cb) | ((cw >> 20) << 16);
    unsigned incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned y = __shfl_up_sync(FULL, incl, d);
      if (lane >= d) incl += y;
    }
    if (j < p) w.rowptr()[j] = row_carry + incl - cnt;
    row_carry += __shfl_sync(FULL, incl, 31);
    const unsigned sd = (cf ? 0x400u : 0u) | (cb ? 0x800u : 0u);
    const unsigned b = __ballot_sync(FULL, sd != 0);
    if (sd) w.act()[n_act + __popc(b & lt)] = ((unsigned)j << 2) | sd;
    n_act += __popc(b);
  }
  if (lane == 0) w.rowptr()[p] = row_carry;
  n_act_out = n_act;

  // stable counting sort of the N list by signal -> csc (cursor holds the counts)
  unsigned carry = 0;
#pragma unroll 1
  for (int i0 = 0; i0 < s; i0 += 32) {
    const int i = i0 + lane;
    const unsigned cnt = i < s ? cursor[i] : 0u;
    unsigned incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned y = __shfl_up_sync(FULL, incl, d);
      if (lane >= d) incl += y;
    }
    if (i < s) {
      w.colptr()[i] = carry + incl - cnt;
      cursor[i] = carry + incl - cnt;
    }
    carry += __shfl_sync(FULL, incl, 31);
  }
  if (lane == 0) w.colptr()[s] = carry;
  __syncwarp();
  unsigned *sorted = w.csc();  // the unsorted list (tmp) is in the nvs half
#pragma unroll 1
  for (int base = 0; base < nnz_tot; base += 32) {
    const int e = base + lane;
    const bool v = e < nnz_tot;
    const unsigned wd = v ? tmp[e] : 0u;
    const unsigned sg = v ? (wd & 0xffu) : (0x100u + lane);
    const unsigned peers = __match_any_sync(FULL, sg);
    if (v) {
      const unsigned pos = cursor[sg] + __popc(peers & lt);
      sorted[pos] = (((wd >> 8) & 0xffu) << 2) | ((unsigned)((int)wd >> 16) << 10);
    }
    __syncwarp();
    if (v && lane == 31 - __clz(peers)) cursor[sg] += __popc(peers);
    __syncwarp();
  }

  // per-protein scalars and the Kmr gather: awaited here, or by the caller before their first use
  if (WAIT_ASYNC) cp_async_wait_all();
  __syncwarp();
  return true;
}

// tuning switches (defaults are the measured best; -D overrides are for the A/B builds of
// profiles/ab_variants.py)
#ifndef MS_TUNE_UNROLL_C
#define MS_TUNE_UNROLL_C 2
#endif
#ifndef MS_TUNE_BOOK_WARP
#define MS_TUNE_BOOK_WARP 7
#endif
constexpr int UNROLL_C = MS_TUNE_UNROLL_C;

// One list entry's factor x^n: same multiplication order as powi(); exponents 1..3 (99.8 %)
// without a branch.  An all-zero word is a neutral entry (factor 1 on the substrate side).
__device__ __forceinline__ float entry_factor(unsigned wd, float x) {
  float r = (wd & 0x400u) ? x : 1.0f;
  float b = x * x;
  const float rb = r * b;
  r = (wd & 0x800u) ? rb : r;
  unsigned n = wd & 0x7ffff000u;  // exponent >> 2, still in place
  if (n) {                        // exponent >= 4
#pragma unroll 1
    do {
      b = b * b;
      if (n & 0x1000u) r = r * b;
      n = (n >> 1) & 0x7ffff000u;
    } while (n);
  }
  return r;
}

// products over a protein's substrate and product entries (kinetics.py:894-918), each side
// in ascending signal order.  xaddr: shared address of X0 or X1 (aligned to its stride).
__device__ __forceinline__ void list_products(unsigned xaddr, unsigned pe, unsigned pend, float &xf,
                                              float &xb) {
  xf = 1.0f;
  xb = 1.0f;
  // W entries per trip; the next trip's entry words are fetched before this trip's signal
  // values, so that one shared-memory latency (not two) separates consecutive trips.  A list
  // whose length is not a multiple of W ends with neutral entries (x * 1.0f is exact).
  constexpr int W = 2;  // 4 was measured slower: short lists pay for the padding
  asm volatile("" : "+r"(xaddr));  // opaque: held in a register instead of being recomputed every trip
  unsigned wd[W];
#pragma unroll
  for (int e = 0; e < W; e++) wd[e] = pe + 4 * e < pend ? lds_u(pe + 4 * e) : 0u;
#pragma unroll 1
  for (; pe < pend; pe += 4 * W) {
    unsigned nx[W];
    float x[W];
#pragma unroll
    for (int e = 0; e < W; e++) nx[e] = pe + 4 * (W + e) < pend ? lds_u(pe + 4 * (W + e)) : 0u;
#pragma unroll
    for (int e = 0; e < W; e++) x[e] = lds_f(xaddr | (wd[e] & 0x3fcu));
#pragma unroll
    for (int e = 0; e < W; e++) {
      const float r = entry_factor(wd[e], x[e]);
      if ((int)wd[e] < 0) xb = xb * r;
      else xf = xf * r;
      wd[e] = nx[e];
    }
  }
  xf = fix_product(xf);
  xb = fix_product(xb);
}

// ---------------------------------------------------------------------------
// equilibrium trimming (kinetics.py:808-859) of ONE cell by one warp, on the cell's lists:
// per iteration the quotients and decisions of the active proteins, then X1 of every signal.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void eq_iterations(const WarpMem &w, int s, int lane, int n_act, int q, int iters,
                                              bool poison, unsigned *sflags) {
  const float qnan = __int_as_float(0x7fc00000);
  const unsigned aX0 = w.aX0(), aX1 = w.aX1(), aV = w.aV(), aF = w.aF();
  const unsigned aRow = w.aRow(), aEnt = w.aEnt(), aCsc = w.aCsc();
  const unsigned dNvs = 4u * (unsigned)(w.oNvs - w.oCsc);  // nvs[e] is at csc[e] + dNvs (mod 2^32)
#pragma unroll 1
  for (int k = 0; k < iters; k++) {
    const float inc = eq_increment(k);
    unsigned mine = 0;
    bool changed = false;
#pragma unroll 1
    for (int t = lane; t < n_act; t += 32) {
      const unsigned aw = lds_u(w.aAct() + 4 * t);
      const unsigned j4 = aw & 0x3fcu;
      const unsigned r0 = lds_u(aRow | j4), r1 = lds_u((aRow | j4) + 4);
      float xf, xb;
      list_products(aX1, aEnt + 4 * (r0 & 0xffffu), aEnt + 4 * (r1 & 0xffffu), xf, xb);
      if (!(aw & 0x400u)) xf = 0.0f;
      if (!(aw & 0x800u)) xb = 0.0f;
      const float F = lds_f(aF | j4);
      const unsigned dec = eq_decide_fast(lds_f(aV | j4), quotient_fixed(xb, xf), lds_f(w.aKe() | j4), F);
      if (dec & 3u) {
        mine |= (dec & 4u) ? 3u : 2u;
        const float Fn = eq_update(F, dec, inc);
        changed |= (Fn != F);
        sts_f(aF | j4, Fn);
      }
    }
    const unsigned any = __reduce_or_sync(FULL, mine);
    // If no F moved (adjustable proteins can sit clamped at 0 or 1), X1 would be recomputed
    // bit-identically and every later iteration of this part repeats this one exactly: same
    // decisions, same flags, no change.  Report the flags for those iterations and stop.
    const bool moved = __any_sync(FULL, changed);
    if (lane == 0 && any) {
      const int k_end = moved ? k + 1 : iters;
      for (int kk = k; kk < k_end; kk++) {
        if (any & 1u) sflags[q * 4 + kk] = 1u;
        sflags[12 + q * 4 + kk] = 1u;
      }
    }
    if (!moved) break;
    __syncwarp();
#pragma unroll 1
    for (int i = lane; i < s; i += 32) {
      const unsigned c0 = aCsc + 4 * lds_u(w.aCol() + 4 * i), c1 = aCsc + 4 * lds_u(w.aCol() + 4 * i + 4);
      float acc = 0.0f;
#pragma unroll 2
      for (unsigned pc = c0; pc < c1; pc += 4)
        acc = acc + lds_f(pc + dNvs) * lds_f(aF | (lds_u(pc) & 0x3fcu));
      float x = clamp0(lds_f(aX0 + 4 * i) + acc);
      if (poison) x = qnan;
      sts_f(aX1 + 4 * i, x);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// sparse path: all parts of one cell on the shared-memory lists
// ---------------------------------------------------------------------------
__device__ void run_sparse(const KinArgs &a, const WarpMem &w, long long cell, int lane, int n_act,
                           unsigned *sflags) {
  const int p = a.p, s = a.s;
  const float qnan = __int_as_float(0x7fc00000);
  const unsigned aX0 = w.aX0(), aX1 = w.aX1(), aV = w.aV(), aF = w.aF(), aFk = w.aFk();
  const unsigned aRow = w.aRow(), aEnt = w.aEnt(), aCsc = w.aCsc();
  const unsigned dNvs = 4u * (unsigned)(w.oNvs - w.oCsc);  // nvs[e] is at csc[e] + dNvs (mod 2^32)
#pragma unroll 1
  for (int q = 0; q < a.n_parts; q++) {
    const float trim = a.trims[q];
    for (int j = lane; j < p; j += 32) {
      sts_f(aV + 4 * j, 0.0f);  // proteins without substrate/product entries have V = 0
      sts_f(aF + 4 * j, 1.0f);
      sts_u(aFk + 4 * j, KEY_ONE);
    }
    __syncwarp();
    // ---- velocities (kinetics.py:771-806) on the active proteins
    bool poison = false;
#pragma unroll 1
    for (int t = lane; t < n_act; t += 32) {
      const unsigned aw = lds_u(w.aAct() + 4 * t);
      const unsigned j4 = aw & 0x3fcu;
      const unsigned r0 = lds_u(aRow | j4), r1 = lds_u((aRow | j4) + 4);
      float xf, xb;
      list_products(aX0, aEnt + 4 * (r0 & 0xffffu), aEnt + 4 * (r1 & 0xffffu), xf, xb);
      float kf = xf / lds_f(w.aKmf() | j4);
      if (!(aw & 0x400u)) kf = 0.0f;
      if (isinf(kf)) kf = V_MAX;
      float kb = xb / lds_f(w.aKmb() | j4);
      if (!(aw & 0x800u)) kb = 0.0f;
      if (isinf(kb)) kb = V_MAX;
      float a_reg = 1.0f;
#pragma unroll 1
      for (unsigned e = r0 >> 16; e < (r1 >> 16); e++) {
        const unsigned wd = lds_u(w.aReg() + 4 * e);
        a_reg = a_reg * reg_factor(lds_f(aX0 + 4 * (wd & 0xffu)), (int)(int8_t)((wd >> 8) & 0xff),
                                   lds_f(w.aRegk() + 4 * e));
      }
      float vmax = lds_f(w.aVmax() | j4);
      if (!a.adj_vmax) vmax = clamp0(vmax * trim);  // kinetics.py:749
      const float v = velocity_from(kf, kb, a_reg, vmax);
      sts_f(aV | j4, v);
      poison |= isnan(v);
    }
    poison = __any_sync(FULL, poison);  // 0 * NaN of the dense reference reaches every signal
    __syncwarp();

    // ---- substrate trimming (kinetics.py:861-879)
#pragma unroll 1
    for (int i = lane; i < s; i += 32) {
      const unsigned c0 = aCsc + 4 * lds_u(w.aCol() + 4 * i), c1 = aCsc + 4 * lds_u(w.aCol() + 4 * i + 4);
      float d = 0.0f;
#pragma unroll 2
      for (unsigned pc = c0; pc < c1; pc += 4) {
        const unsigned wd = lds_u(pc);
        const float nv = (float)((int)wd >> 10) * lds_f(aV | (wd & 0x3fcu));  // kinetics.py:759
        sts_f(pc + dNvs, nv);
        float neg = -nv;
        neg = neg < 0.0f ? 0.0f : neg;
        d = d + neg;
      }
      float f = lds_f(aX0 + 4 * i) / d;
      if (f > 1.0f) f = 1.0f;
      const unsigned key = fkey(f);
#pragma unroll 2
      for (unsigned pc = c0; pc < c1; pc += 4)  // every protein that removes signal i is limited by F_i
        if (lds_f(pc + dNvs) < 0.0f) atom_min_shared(aFk | (lds_u(pc) & 0x3fcu), key);
    }
    __syncwarp();
    // X1 = max(X0 + sum_j NV'_ji, 0)   (kinetics.py:763-764)
#pragma unroll 1
    for (int i = lane; i < s; i += 32) {
      const unsigned c0 = aCsc + 4 * lds_u(w.aCol() + 4 * i), c1 = aCsc + 4 * lds_u(w.aCol() + 4 * i + 4);
      float acc = 0.0f;
#pragma unroll 2
      for (unsigned pc = c0; pc < c1; pc += 4) {
        const float nvp = lds_f(pc + dNvs) * fkey_decode(lds_u(aFk | (lds_u(pc) & 0x3fcu)));
        sts_f(pc + dNvs, nvp);
        acc = acc + nvp;
      }
      float x = clamp0(lds_f(aX0 + 4 * i) + acc);
      if (poison) x = qnan;
      sts_f(aX1 + 4 * i, x);
    }
    __syncwarp();

    // ---- equilibrium trimming (kinetics.py:808-859)
    eq_iterations(w, s, lane, n_act, q, a.eq_iters[q], poison, sflags);

    if (a.V_out)
      for (int j = lane; j < p; j += 32) a.V_out[cell * (long long)p + j] = lds_f(aV + 4 * j);
    // next part starts from this part's result
    for (int i = lane; i < s; i += 32) sts_f(aX0 + 4 * i, lds_f(aX1 + 4 * i));
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// batch path: the 8 warps of a CTA integrate their 8 cells TOGETHER.  In the warp-per-cell
// path a sweep costs as many loop iterations as the longest list among 32 proteins (signals),
// three times the average.  Here the proteins (and the signals) of all 8 cells are pooled and
// sorted by list length, so the 32 lanes of a warp walk lists of equal length.  Every item is
// computed exactly as in run_sparse (same loops, same order), so results are bit-identical;
// only the order in which independent items are visited changes.
// ---------------------------------------------------------------------------
constexpr int BATCH = 8;            // cells per CTA = warps per CTA
constexpr int BATCH_MAX_PS = 64;    // p and s limit of the batch path (item tables below)
constexpr unsigned CS_IMP = 2u, CS_ANY = 4u, CS_MOVED = 8u;

constexpr int ITEMS_PER_WARP = BATCH_MAX_PS;  // 8 * 64 pooled items over 8 warps

// The sorted item sequence is dealt to the warps in chunks of 32, and each warp's chunks are stored
// contiguously: sorted position -> index in pitem / citem.  A sweep lasts as long as its slowest
// warp (the sweeps are separated by CTA barriers), and that is the warp with chunk 0, the longest
// lists.  The chunks after the first eight are therefore dealt backwards (chunk 8 to warp 7, ...),
// and the per-iteration flag bookkeeping is done by lanes of warp 7, which has the shortest lists.
__device__ __forceinline__ unsigned item_index(unsigned pos) {
  const unsigned chunk = pos >> 5;
  const unsigned owner = chunk < BATCH ? chunk : 2 * BATCH - 1 - chunk;
  return owner * ITEMS_PER_WARP + (chunk / BATCH) * 32 + (pos & 31u);
}
// number of the n sorted items that land in `warp`'s list
__device__ __forceinline__ int items_of_warp(int n, int warp) {
  const int a = n - 32 * warp, b = n - 32 * (2 * BATCH - 1 - warp);
  return min(max(a, 0), 32) + min(max(b, 0), 32);
}

struct BatchShared {
  unsigned short pitem[BATCH * ITEMS_PER_WARP];  // slot << 12 | act word (protein*4 | side flags)
  unsigned short citem[BATCH * ITEMS_PER_WARP];  // slot << 8 | signal
  unsigned cnt_p[17], cnt_c[33];                 // items per list-length bucket (longest lists first)
  unsigned cur_p[17], cur_c[33];                 // placement cursors inside the buckets
  unsigned cell_state[2][BATCH];                 // CS_IMP/ANY/MOVED of equilibrium iteration k in [k & 1]
  unsigned cell_poison[BATCH];                   // a velocity of the cell is NaN in this part
  long long next_base;
};

// Regulatory factors of one cell from its X0 (kinetics.py:789-799), one lane per entry, into the
// cell's X1 + cursor words (2*S2 >= e_reg; X1 is only written after the velocity sweep, the cursor
// only while the lists are built); the velocity sweep multiplies them up in list order.  They depend on X0 only, so every warp does this for its own cell outside the
// sweeps, where the 1-3 entries of a regulated protein would otherwise prolong its lane.
__device__ __forceinline__ void regulation_pass(unsigned sb, unsigned oRow, unsigned oReg, unsigned oRegk,
                                                unsigned oFac, int p, int lane) {
  const unsigned n_reg = lds_u(sb + oRow + 4u * (unsigned)p) >> 16;
  for (unsigned e = lane; e < n_reg; e += 32) {
    const unsigned wd = lds_u(sb + oReg + 4 * e);
    sts_f(sb + oFac + 4 * e, reg_factor(lds_f(sb + 4 * (wd & 0xffu)), (int)(int8_t)((wd >> 8) & 0xff),
                                        lds_f(sb + oRegk + 4 * e)));
  }
}

// On entry (and again at the end of every part) V = 0, F = 1, FminKey = 1 in every present cell's
// region, its regulatory factors are in place, cell_state and cell_poison are zero, and a CTA
// barrier has been passed.
__device__ void run_batch(const KinArgs &a, const WarpMem &w0, unsigned region_bytes, BatchShared &bs,
                          const long long *cells /* per slot, -1 = absent */, int n_pitems, int n_citems,
                          unsigned *sflags) {
  int tid = threadIdx.x;
  asm volatile("mov.b32 %0, %0;" : "+r"(tid));  // opaque: kept, not re-read from the special register in the loops
  const int lane = tid & 31, slot_own = tid >> 5;
  const int p = a.p, s = a.s;
  const float qnan = __int_as_float(0x7fc00000);
  const unsigned dNvs = 4u * (unsigned)(w0.oNvs - w0.oCsc);
  const bool own = cells[slot_own] >= 0;
  const unsigned sb_own = w0.sb + (unsigned)slot_own * region_bytes;
  const unsigned oX1 = w0.aX1() - w0.sb, oV = w0.aV() - w0.sb, oF = w0.aF() - w0.sb, oFk = w0.aFk() - w0.sb;
  const unsigned oRow = w0.aRow() - w0.sb, oEnt = w0.aEnt() - w0.sb, oCsc = w0.aCsc() - w0.sb;
  const unsigned oCol = w0.aCol() - w0.sb, oKe = w0.aKe() - w0.sb, oKmf = w0.aKmf() - w0.sb;
  const unsigned oKmb = w0.aKmb() - w0.sb, oVmax = w0.aVmax() - w0.sb, oReg = w0.aReg() - w0.sb;
  const unsigned oRegk = w0.aRegk() - w0.sb, oAct = w0.aAct() - w0.sb;
  // this warp's share of the pooled items
  // (32-bit shared-window addresses, opaque so that they are held, not rebuilt from the CTA's
  // shared-window base at every access; aState: cell_state[2][8] followed by cell_poison[8])
  unsigned aPmine = (unsigned)__cvta_generic_to_shared(bs.pitem + slot_own * ITEMS_PER_WARP);
  unsigned aCmine = (unsigned)__cvta_generic_to_shared(bs.citem + slot_own * ITEMS_PER_WARP);
  unsigned aState = (unsigned)__cvta_generic_to_shared(&bs.cell_state[0][0]);
  asm volatile("" : "+r"(aPmine), "+r"(aCmine), "+r"(aState));
  const unsigned aPoison = aState + 8u * BATCH;
  const int n_p = items_of_warp(n_pitems, slot_own), n_c = items_of_warp(n_citems, slot_own);
  if (a.n_parts <= 0) __syncthreads();  // callers rely on a barrier at the end

#pragma unroll 1
  for (int q = 0; q < a.n_parts; q++) {
    const float trim = a.trims[q];
    const int iters = a.eq_iters[q];
    // ---- velocities (kinetics.py:771-806)
#pragma unroll 1
    for (int it = lane; it < n_p; it += 32) {
      const unsigned item = lds_u16(aPmine + 2 * it);
      unsigned sb = w0.sb + (item >> 12) * region_bytes;
      asm volatile("" : "+r"(sb));  // held, not recomputed at each use
      const unsigned aw = item & 0xfffu, j4 = aw & 0x3fcu;
      const unsigned r0 = lds_u((sb + oRow) | j4), r1 = lds_u(((sb + oRow) | j4) + 4);
      float xf, xb;
      list_products(sb, sb + oEnt + 4 * (r0 & 0xffffu), sb + oEnt + 4 * (r1 & 0xffffu), xf, xb);
      float kf = xf / lds_f((sb + oKmf) | j4);
      if (!(aw & 0x400u)) kf = 0.0f;
      if (isinf(kf)) kf = V_MAX;
      float kb = xb / lds_f((sb + oKmb) | j4);
      if (!(aw & 0x800u)) kb = 0.0f;
      if (isinf(kb)) kb = V_MAX;
      float a_reg = 1.0f;
#pragma unroll 1
      for (unsigned e = r0 >> 16; e < (r1 >> 16); e++) a_reg = a_reg * lds_f(sb + oX1 + 4 * e);
      float vmax = lds_f((sb + oVmax) | j4);
      if (!a.adj_vmax) vmax = clamp0(vmax * trim);
      const float v = velocity_from(kf, kb, a_reg, vmax);
      sts_f((sb + oV) | j4, v);
      if (isnan(v)) sts_u(aPoison + 4 * (item >> 12), 1u);
    }
    __syncthreads();
    // ---- substrate trimming (kinetics.py:861-879)
#pragma unroll 1
    for (int it = lane; it < n_c; it += 32) {
      const unsigned item = lds_u16(aCmine + 2 * it);
      unsigned sb = w0.sb + (item >> 8) * region_bytes;
      asm volatile("" : "+r"(sb));  // held, not recomputed at each use
      const unsigned i = item & 0xffu;
      const unsigned c0 = sb + oCsc + 4 * lds_u(sb + oCol + 4 * i), c1 = sb + oCsc + 4 * lds_u(sb + oCol + 4 * i + 4);
      const unsigned aV = sb + oV, aFk = sb + oFk;
      float d = 0.0f;
#pragma unroll UNROLL_C
      for (unsigned pc = c0, pn = c0 + dNvs; pc < c1; pc += 4, pn += 4) {
        const unsigned wd = lds_u(pc);
        const float nv = (float)((int)wd >> 10) * lds_f(aV | (wd & 0x3fcu));
        sts_f(pn, nv);
        float neg = -nv;
        neg = neg < 0.0f ? 0.0f : neg;
        d = d + neg;
      }
      float f = lds_f(sb + 4 * i) / d;
      if (f > 1.0f) f = 1.0f;
      const unsigned key = fkey(f);
#pragma unroll UNROLL_C
      for (unsigned pc = c0, pn = c0 + dNvs; pc < c1; pc += 4, pn += 4)
        if (lds_f(pn) < 0.0f) atom_min_shared(aFk | (lds_u(pc) & 0x3fcu), key);
    }
    __syncthreads();
    // X1 = max(X0 + sum_j NV'_ji, 0)
#pragma unroll 1
    for (int it = lane; it < n_c; it += 32) {
      const unsigned item = lds_u16(aCmine + 2 * it);
      unsigned slot = item >> 8, sb = w0.sb + slot * region_bytes, i = item & 0xffu;
      asm volatile("" : "+r"(sb));  // held, not recomputed at each use
      const unsigned c0 = sb + oCsc + 4 * lds_u(sb + oCol + 4 * i), c1 = sb + oCsc + 4 * lds_u(sb + oCol + 4 * i + 4);
      const unsigned aFk = sb + oFk;
      float acc = 0.0f;
#pragma unroll UNROLL_C
      for (unsigned pc = c0, pn = c0 + dNvs; pc < c1; pc += 4, pn += 4) {
        const float nvp = lds_f(pn) * fkey_decode(lds_u(aFk | (lds_u(pc) & 0x3fcu)));
        sts_f(pn, nvp);
        acc = acc + nvp;
      }
      float x = clamp0(lds_f(sb + 4 * i) + acc);
      if (lds_u(aPoison + 4 * slot)) x = qnan;
      sts_f(sb + oX1 + 4 * i, x);
    }
    __syncthreads();

    // ---- equilibrium trimming (kinetics.py:808-859)
    // A cell iterates as long as one of its F moved in the previous iteration (see run_sparse);
    // CS_MOVED of iteration k - 1 is therefore the cell's "still iterating" flag of iteration k.
#pragma unroll 1
    for (int k = 0; k < iters; k++) {
      const float inc = eq_increment(k);
      unsigned *st_cur = bs.cell_state[k & 1], *st_prv = bs.cell_state[(k & 1) ^ 1];
      const unsigned a_cur = aState + 4u * BATCH * (unsigned)(k & 1), a_prv = aState + 4u * BATCH * (unsigned)((k & 1) ^ 1);
      int moved_mine = 0;
#pragma unroll 1
      for (int it = lane; it < n_p; it += 32) {
        const unsigned item = lds_u16(aPmine + 2 * it);
        const unsigned slot = item >> 12;
        if (k > 0 && !(lds_u(a_prv + 4 * slot) & CS_MOVED)) continue;
        unsigned sb = w0.sb + slot * region_bytes;
        asm volatile("" : "+r"(sb));  // held, not recomputed at each use
        const unsigned aw = item & 0xfffu, j4 = aw & 0x3fcu;
        const unsigned r0 = lds_u((sb + oRow) | j4), r1 = lds_u(((sb + oRow) | j4) + 4);
        float xf, xb;
        list_products(sb + oX1, sb + oEnt + 4 * (r0 & 0xffffu), sb + oEnt + 4 * (r1 & 0xffffu), xf, xb);
        if (!(aw & 0x400u)) xf = 0.0f;
        if (!(aw & 0x800u)) xb = 0.0f;
        const float F = lds_f((sb + oF) | j4);
        const unsigned dec = eq_decide_fast(lds_f((sb + oV) | j4), quotient_fixed(xb, xf), lds_f((sb + oKe) | j4), F);
        if (dec & 3u) {
          const float Fn = eq_update(F, dec, inc);
          sts_f((sb + oF) | j4, Fn);
          red_or_shared(a_cur + 4 * slot, ((dec & 4u) ? (CS_IMP | CS_ANY) : CS_ANY) | (Fn != F ? CS_MOVED : 0u));
          moved_mine |= (Fn != F) ? 1 : 0;
        }
      }
      const int any_moved = __syncthreads_or(moved_mine);
      if (slot_own == MS_TUNE_BOOK_WARP && lane < BATCH) {  // flags of this iteration; runs beside the X1 sweep
        const int tid = lane;
        const unsigned st = st_cur[tid];
        if (st & CS_ANY) {  // see run_sparse: an unmoved cell repeats this iteration exactly
          const int k_end = (st & CS_MOVED) ? k + 1 : iters;
          for (int kk = k; kk < k_end; kk++) {
            if (st & CS_IMP) sflags[q * 4 + kk] = 1u;
            sflags[12 + q * 4 + kk] = 1u;
          }
        }
        st_prv[tid] = 0;  // becomes iteration k + 1's state; its readers finished before the barrier
      }
      if (!any_moved) break;
#pragma unroll 1
      for (int it = lane; it < n_c; it += 32) {
        const unsigned item = lds_u16(aCmine + 2 * it);
        const unsigned slot = item >> 8;
        if (!(lds_u(a_cur + 4 * slot) & CS_MOVED)) continue;
        unsigned sb = w0.sb + slot * region_bytes, i = item & 0xffu;
        asm volatile("" : "+r"(sb));  // held, not recomputed at each use
        const unsigned c0 = sb + oCsc + 4 * lds_u(sb + oCol + 4 * i), c1 = sb + oCsc + 4 * lds_u(sb + oCol + 4 * i + 4);
        const unsigned aF = sb + oF;
        float acc = 0.0f;
#pragma unroll UNROLL_C
        for (unsigned pc = c0, pn = c0 + dNvs; pc < c1; pc += 4, pn += 4)
          acc = acc + lds_f(pn) * lds_f(aF | (lds_u(pc) & 0x3fcu));
        float x = clamp0(lds_f(sb + 4 * i) + acc);
        if (lds_u(aPoison + 4 * slot)) x = qnan;
        sts_f(sb + oX1 + 4 * i, x);
      }
      __syncthreads();
    }

    // this part's result is the next part's start; arrays and flags back to their entry state
    if (own) {
      if (a.V_out)
        for (int j = lane; j < p; j += 32) a.V_out[cells[slot_own] * (long long)p + j] = lds_f(sb_own + oV + 4 * j);
      for (int i = lane; i < s; i += 32) sts_f(sb_own + 4 * i, lds_f(sb_own + oX1 + 4 * i));
      __syncwarp();
      if (q + 1 < a.n_parts) regulation_pass(sb_own, oRow, oReg, oRegk, oX1, p, lane);
      for (int j = lane; j < p; j += 32) {
        sts_f(sb_own + oV + 4 * j, 0.0f);  // proteins without substrate/product entries have V = 0
        sts_f(sb_own + oF + 4 * j, 1.0f);
        sts_u(sb_own + oFk + 4 * j, KEY_ONE);
      }
    }
    // Reset by the lanes that read the states for the flags above: on the early `break` no barrier
    // separates that read from this write, program order does.  (cell_poison was last read in an
    // X1 sweep, before the loop's last barrier.)
    if (slot_own == MS_TUNE_BOOK_WARP && lane < BATCH) {
      bs.cell_state[0][lane] = 0;
      bs.cell_state[1][lane] = 0;
      bs.cell_poison[lane] = 0;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// dense path: same arithmetic straight from global memory (any p, s, any int32)
// F, V, Fmin live in `scr` (global, 3*p floats per warp) when p is too large for
// the shared arrays; X0/X1/Fsig always fit (s32 words are reserved per warp).
// ---------------------------------------------------------------------------
__device__ float dense_product(const float *X, const int *K, int s, bool &involved) {
  float xx = 1.0f;
  bool any = false;
  for (int i = 0; i < s; i++) {
    const int k = K[i];
    if (k > 0) {
      any = true;
      xx = xx * powi(X[i], k);
    } else if (k < 0) {
      xx = xx * __int_as_float(0x7f800000);
    }
  }
  involved = any;
  return fix_product(xx);
}

__device__ __noinline__ void run_dense(const KinArgs &a, float *X0, float *X1, float *Fsig, float *V, float *Fmin,
                          float *F, long long cell, int lane, unsigned *sflags) {
  const int p = a.p, s = a.s;
  // adj_vmax / V_out (logical cell index) are never combined with a row map (launch_fused)
  const long long row = param_row(a, cell) * (long long)p;
  const long long tile = row * s;
  const int *gN = a.N + tile, *gNf = a.Nf + tile, *gNb = a.Nb + tile, *gA = a.A + tile;
  const float *gKmr = a.Kmr + tile;
  for (int q = 0; q < a.n_parts; q++) {
    const float trim = a.trims[q];
    bool poison = false;
    for (int j = lane; j < p; j += 32) {
      bool inv;
      float kf = dense_product(X0, gNf + j * s, s, inv) / a.Kmf[row + j];
      if (!inv) kf = 0.0f;
      if (isinf(kf)) kf = V_MAX;
      float kb = dense_product(X0, gNb + j * s, s, inv) / a.Kmb[row + j];
      if (!inv) kb = 0.0f;
      if (isinf(kb)) kb = V_MAX;
      float a_reg = 1.0f;
      for (int i = 0; i < s; i++) {
        const int av = gA[j * s + i];
        if (av != 0) a_reg = a_reg * reg_factor(X0[i], av, gKmr[j * s + i]);
      }
      float vmax = a.adj_vmax ? a.adj_vmax[row + j] : clamp0(a.Vmax[row + j] * trim);
      const float v = velocity_from(kf, kb, a_reg, vmax);
      V[j] = v;
      poison |= isnan(v);
    }
    poison = __any_sync(FULL, poison);
    __syncwarp();
    for (int i = lane; i < s; i += 32) {
      float d = 0.0f;
      for (int j = 0; j < p; j++) {
        float neg = -((float)gN[j * s + i] * V[j]);
        neg = neg < 0.0f ? 0.0f : neg;
        d = d + neg;
      }
      float f = X0[i] / d;
      if (f > 1.0f) f = 1.0f;
      Fsig[i] = f;
    }
    __syncwarp();
    for (int j = lane; j < p; j += 32) {
      const float v = V[j];
      float fm = 1.0f;
      for (int i = 0; i < s; i++)
        if ((float)gN[j * s + i] * v < 0.0f) fm = nanmin(fm, Fsig[i]);
      Fmin[j] = fm;
      F[j] = 1.0f;
    }
    __syncwarp();
    for (int i = lane; i < s; i += 32) {
      float acc = 0.0f;
      for (int j = 0; j < p; j++) acc = acc + ((float)gN[j * s + i] * V[j]) * Fmin[j];
      X1[i] = clamp0(X0[i] + acc);
    }
    __syncwarp();
    const int iters = a.eq_iters[q];
    for (int k = 0; k < iters; k++) {
      const float inc = eq_increment(k);
      unsigned mine = 0;
      for (int j = lane; j < p; j += 32) {
        bool invb, invf;
        float xb = dense_product(X1, gNb + j * s, s, invb);
        if (!invb) xb = 0.0f;
        float xf = dense_product(X1, gNf + j * s, s, invf);
        if (!invf) xf = 0.0f;
        const float Fj = F[j];
        const unsigned dec = eq_decide(V[j], quotient_from(xb, xf), a.Ke[row + j], Fj);
        if (dec & 3u) {
          mine |= (dec & 4u) ? 3u : 2u;
          F[j] = eq_update(Fj, dec, inc);
        }
      }
      const unsigned any = __reduce_or_sync(FULL, mine);
      if (lane == 0 && any) {
        if (any & 1u) sflags[q * 4 + k] = 1u;
        sflags[12 + q * 4 + k] = 1u;
      }
      if (!any) break;
      __syncwarp();
      for (int i = lane; i < s; i += 32) {
        float acc = 0.0f;
        for (int j = 0; j < p; j++)
          acc = acc + (((float)gN[j * s + i] * V[j]) * Fmin[j]) * F[j];
        X1[i] = clamp0(X0[i] + acc);
      }
      __syncwarp();
    }
    (void)poison;  // dense sums include 0 * NaN terms themselves
    if (a.V_out)
      for (int j = lane; j < p; j += 32) a.V_out[row + j] = V[j];
    for (int i = lane; i < s; i += 32) X0[i] = X1[i];
    __syncwarp();
  }
}

}  // namespace ms
#include "kinetics_lanes.cuh"
namespace ms {

// ---------------------------------------------------------------------------
// The lane path as its own kernel: every warp takes cells from the work counter and integrates them
// alone (kinetics_lanes.cuh).  Kept apart from kinetics_fused_kernel so that the code the 32
// independent warps of an SM run through (scan, list building, integration) stays small: with the
// warps at different places of it, instruction fetch decides the speed.  A cell that does not fit
// the fixed-format lists is appended to the deferred list (tier 2 of kinetics_fused_kernel).
// ---------------------------------------------------------------------------
// X0 of a cell: read, or gathered from cell_molecules and the cell's pixel (world.py:619-621); lane i
// holds signals i (.x) and i + 32 (.y).  One copy of the code for both call sites; everything is
// passed by value (a reference to the kernel parameters would turn into generic loads and a stack
// frame in a function that is not inlined).
__device__ __noinline__ float4 load_cell_signals(const float *X_in, const float *cell_mols, const float *mol_map,
                                                 const int *pos_all, int s, int m, int H, int W, long long cell,
                                                 int lane) {
  float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);  // .z != 0: the cell's position is outside the map
  if (X_in) {
    const float *x = X_in + cell * s;
    if (lane < s) v.x = x[lane];
    if (lane + 32 < s) v.y = x[lane + 32];
    return v;
  }
  // lane i < m: intracellular molecule i; lane m + k: the cell's pixel of map plane k
  int2 pos = *reinterpret_cast<const int2 *>(pos_all + 2 * cell);
  if ((unsigned)pos.x >= (unsigned)H || (unsigned)pos.y >= (unsigned)W) {
    pos = make_int2(0, 0);  // never read outside the map; reported through MS_FLAG_BAD_POSITION
    v.z = 1.0f;
  }
  const unsigned plane = (unsigned)H * (unsigned)W;  // < 2^31 pixels per plane (ms_enzymatic_activity checks)
  const float *px = mol_map + (size_t)pos.x * W + pos.y;
  const float *cm = cell_mols + cell * m;
  if (lane < s) v.x = lane < m ? cm[lane] : px[(size_t)(unsigned)(lane - m) * plane];
  if (lane + 32 < s) v.y = lane + 32 < m ? cm[lane + 32] : px[(size_t)(unsigned)(lane + 32 - m) * plane];
  return v;
}

template <int VEC>
__global__ void __launch_bounds__(MAX_THREADS, MS_TUNE_MIN_CTAS) kinetics_lanes_kernel(const __grid_constant__ KinArgs a) {
  extern __shared__ __align__(16) unsigned smem[];
  __shared__ unsigned sflags[MS_N_FLAGS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < MS_N_FLAGS) sflags[threadIdx.x] = 0;
  __syncthreads();
  const unsigned raw = (unsigned)__cvta_generic_to_shared(smem);
  unsigned aligned = (raw + (unsigned)a.smem_align - 1u) & ~((unsigned)a.smem_align - 1u);
  asm volatile("mov.b32 %0, %0;" : "+r"(aligned));  // held in a register, not recomputed at every use
  unsigned *b = smem + ((aligned - raw) >> 2) + (size_t)warp * a.words_per_warp;
  const unsigned sb = aligned + (unsigned)warp * (unsigned)a.words_per_warp * 4u;
  const int p = a.p, s = a.s;
  const long long n_items = a.c;
  const LaneLayout &L = a.lanes;
  auto grab = [&]() -> long long {
    int v = 0;
    if (lane == 0) v = atomicAdd(a.work_counter, 1);
    return (long long)__shfl_sync(FULL, v, 0);
  };
  float xa = 0.0f, xb = 0.0f;
  unsigned wflags = 0;  // flags of this warp's cells (bit = index in the flags buffer)
  long long cell = grab();
  if (cell < n_items) {
    const float4 v = load_cell_signals(a.X_in, a.cell_mols, a.mol_map, a.pos, s, a.m, a.H, a.W, cell, lane);
    xa = v.x;
    xb = v.y;
    if (v.z != 0.0f) wflags |= 1u << MS_FLAG_BAD_POSITION;
  }
#pragma unroll 1
  while (cell < n_items) {
    const long long next_cell = grab();
    const long long prm = param_row(a, cell);
    unsigned *rec = a.cache ? a.cache + prm * (long long)a.rec_words : nullptr;
    const unsigned hdr = rec ? __ldg(rec) : 0u;
    const unsigned status = hdr >> 24;
    LaneCounts cnt = {0, 0, 0};
    bool ok = false;
    if (status == REC_VALID) {
      // a valid record holds exactly what build_lane_lists left in the static arrays
      const uint4 *src = reinterpret_cast<const uint4 *>(rec + REC_HEADER_WORDS);
      const int n16 = L.static_words >> 2;
#pragma unroll 1
      for (int e = lane; e < n16; e += 32) cp_async16(sb + 16u * (unsigned)e, src + e, 0u);
      cnt = unpack_counts(hdr);
      ok = true;
    } else if (status != REC_OVERFLOW) {
      const long long tile = prm * (long long)p * s;
      int nq;
      if (VEC == 4)
        nq = scan_tile_lanes(a.Nf + tile, a.Nb + tile, a.N + tile, a.A + tile, p, s, sb,
                             sb + 4u * (unsigned)L.oQueue, L.q_cap, lane);
      else
        nq = scan_tile<VEC>(a, b + L.oQueue, L.q_cap, tile, lane);
      __syncwarp();
      ok = nq >= 0 && build_lane_lists(a, L, b, sb, prm, lane, nq, cnt);
      if (rec) {  // keep the lists for the next steps
        if (ok) {
          cp_async_wait_all();  // scalars and Kmr have landed
          __syncwarp();
          uint4 *dst = reinterpret_cast<uint4 *>(rec + REC_HEADER_WORDS);
          const uint4 *seg = reinterpret_cast<const uint4 *>(b);
          const int n16 = L.static_words >> 2;
#pragma unroll 1
          for (int e = lane; e < n16; e += 32) dst[e] = seg[e];
        }
        if (lane == 0) rec[0] = ok ? (REC_VALID << 24) | pack_counts(cnt) : REC_OVERFLOW << 24;
      }
    }
    if (ok) {
      if (lane < s) sts_f(sb + 4u * (unsigned)(L.oX0 + lane), xa);
      if (lane + 32 < s) sts_f(sb + 4u * (unsigned)(L.oX0 + lane + 32), xb);
    }
    if (next_cell < n_items) {  // in flight during this cell's parts
      const float4 v = load_cell_signals(a.X_in, a.cell_mols, a.mol_map, a.pos, s, a.m, a.H, a.W, next_cell, lane);
      xa = v.x;
      xb = v.y;
      if (v.z != 0.0f) wflags |= 1u << MS_FLAG_BAD_POSITION;
    }
    cp_async_wait_all();
    __syncwarp();
    if (ok) {
      run_lanes(a, L, sb, cell, lane, cnt, wflags);
#pragma unroll 1
      for (int i = lane; i < s; i += 32) a.X_out[cell * s + i] = lds_f(sb + 4u * (unsigned)(L.oX0 + i));
    } else if (lane == 0) {
      a.defer_list[atomicAdd(a.defer_count, 1)] = (int)cell;  // integrated by the tier-2 launch
    }
    __syncwarp();
    cell = next_cell;
  }
  if (lane < MS_N_FLAGS && ((wflags >> lane) & 1u)) sflags[lane] = 1u;
  __syncthreads();
  if (threadIdx.x < MS_N_FLAGS && sflags[threadIdx.x]) a.flags[threadIdx.x] = 1u;
}

// ---------------------------------------------------------------------------
// The lane path with the dense tiles read in every launch (no list cache), phased per SM.
// In kinetics_lanes_kernel a warp scans, builds and integrates its cell in one go; with 32 warps of an
// SM at different places of those 46 KB of code the instruction cache (32 KB) misses all the time
// (28 % of the warp samples wait for instructions).  Here ONE CTA of 32 warps owns the SM and takes
// rounds of `phase_batch` cells: first all warps scan tiles and build lists (records go to a scratch
// slot in global memory, i.e. L2), then, after a CTA barrier, all warps integrate from the records.
// Each phase's code fits the cache; cells are handed out dynamically inside a phase, so a round
// ends at most one cell's work after its average warp.  Same device functions, same lists, same
// bits as kinetics_lanes_kernel.
// ---------------------------------------------------------------------------
constexpr int PHASED_THREADS = 1024;

// the four dense (p,s) int32 tiles of a parameter row on their way into L2 (this GPU's memory), one
// bulk prefetch each: the scan that follows a cell-time later finds them there instead of in HBM
__device__ __forceinline__ void prefetch_tiles(const KinArgs &a, long long cell, int lane) {
  if (lane >= 4) return;
  const long long tile = param_row(a, cell) * (long long)a.p * a.s;
  const int *src = (lane == 0 ? a.Nf : lane == 1 ? a.Nb : lane == 2 ? a.N : a.A) + tile;
  const unsigned bytes = (unsigned)a.p * (unsigned)a.s * 4u;  // a multiple of 16 (VEC == 4: s % 4 == 0)
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

template <int VEC>
__global__ void __launch_bounds__(PHASED_THREADS, 1) kinetics_lanes_phased_kernel(const __grid_constant__ KinArgs a) {
  extern __shared__ __align__(16) unsigned smem[];
  __shared__ unsigned sflags[MS_N_FLAGS];
  __shared__ long long s_base_next;
  __shared__ int s_next[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < MS_N_FLAGS) sflags[threadIdx.x] = 0;
  const unsigned raw = (unsigned)__cvta_generic_to_shared(smem);
  unsigned aligned = (raw + (unsigned)a.smem_align - 1u) & ~((unsigned)a.smem_align - 1u);
  asm volatile("mov.b32 %0, %0;" : "+r"(aligned));
  unsigned *b = smem + ((aligned - raw) >> 2) + (size_t)warp * a.words_per_warp;
  const unsigned sb = aligned + (unsigned)warp * (unsigned)a.words_per_warp * 4u;
  const int p = a.p, s = a.s;
  const long long n_items = a.c;
  const LaneLayout &L = a.lanes;
  const int rec_words = REC_HEADER_WORDS + L.static_words;
  unsigned *slots = a.phase_recs + (size_t)blockIdx.x * (size_t)a.phase_batch * (size_t)rec_words;
  const int n16 = L.static_words >> 2;
  const int AHEAD = a.phase_ahead;  // a warp that takes item i sends item i + AHEAD on its way
  unsigned wflags = 0;
  auto take = [&](int phase) -> int {
    int v = 0;
    if (lane == 0) v = atomicAdd(&s_next[phase], 1);
    return __shfl_sync(FULL, v, 0);
  };
  if (threadIdx.x == 0) s_base_next = (long long)atomicAdd(a.work_counter, a.phase_batch);
  __syncthreads();
  {  // the first round starts cold: at least its first items are requested before anything else
    const long long first = s_base_next + warp;
    if (warp < AHEAD && first < n_items) prefetch_tiles(a, first, lane);
  }
#pragma unroll 1
  for (;;) {
    const long long base = s_base_next;
    __syncthreads();  // everyone holds this round's base; the round before is done (records, counters)
    if (base >= n_items) break;
    if (threadIdx.x == 0) {
      s_next[0] = 0;
      s_next[1] = 0;
      s_base_next = (long long)atomicAdd(a.work_counter, a.phase_batch);  // the round after this one
    }
    __syncthreads();
    const int n = (int)(n_items - base < (long long)a.phase_batch ? n_items - base : (long long)a.phase_batch);

    // ---- phase 1: tiles -> lists -> record slots
#pragma unroll 1
    for (int i = take(0); i < n; i = take(0)) {
      if (AHEAD > 0 && i + AHEAD < n) prefetch_tiles(a, base + i + AHEAD, lane);
      const long long prm = param_row(a, base + i);
      const long long tile = prm * (long long)p * s;
      LaneCounts cnt = {0, 0, 0};
      const int nq = scan_tile_lanes(a.Nf + tile, a.Nb + tile, a.N + tile, a.A + tile, p, s, sb,
                                     sb + 4u * (unsigned)L.oQueue, L.q_cap, lane);
      __syncwarp();
      const bool ok = nq >= 0 && build_lane_lists(a, L, b, sb, prm, lane, nq, cnt);
      unsigned *rec = slots + (size_t)i * rec_words;
      if (ok) {
        cp_async_wait_all();  // scalars and Kmr have landed
        __syncwarp();
        uint4 *dst = reinterpret_cast<uint4 *>(rec + REC_HEADER_WORDS);
        const uint4 *seg = reinterpret_cast<const uint4 *>(b);
#pragma unroll 1
        for (int e = lane; e < n16; e += 32) dst[e] = seg[e];
      }
      if (lane == 0) rec[0] = ok ? (REC_VALID << 24) | pack_counts(cnt) : REC_OVERFLOW << 24;
      __syncwarp();
    }
    __syncthreads();  // all records of the round are written (and visible to the CTA)
    {  // the next round's first items travel to L2 while this round is integrated
      const long long ahead = s_base_next + warp;
      if (warp < AHEAD && ahead < n_items) prefetch_tiles(a, ahead, lane);
    }

    // ---- phase 2: records -> three parts -> X_out; the next cell's signals are fetched ahead
    int i = take(1);
    float xa = 0.0f, xb = 0.0f;
    if (i < n) {
      const float4 v = load_cell_signals(a.X_in, a.cell_mols, a.mol_map, a.pos, s, a.m, a.H, a.W, base + i, lane);
      xa = v.x;
      xb = v.y;
      if (v.z != 0.0f) wflags |= 1u << MS_FLAG_BAD_POSITION;
    }
#pragma unroll 1
    while (i < n) {
      const int i_next = take(1);
      const long long cell = base + i;
      const unsigned *rec = slots + (size_t)i * rec_words;
      const unsigned hdr = *reinterpret_cast<const volatile unsigned *>(rec);
      const bool ok = (hdr >> 24) == REC_VALID;
      LaneCounts cnt = {0, 0, 0};
      if (ok) {
        const uint4 *src = reinterpret_cast<const uint4 *>(rec + REC_HEADER_WORDS);
#pragma unroll 1
        for (int e = lane; e < n16; e += 32) cp_async16(sb + 16u * (unsigned)e, src + e, 0u);
        cnt = unpack_counts(hdr);
        if (lane < s) sts_f(sb + 4u * (unsigned)(L.oX0 + lane), xa);
        if (lane + 32 < s) sts_f(sb + 4u * (unsigned)(L.oX0 + lane + 32), xb);
      }
      if (i_next < n) {
        const float4 v = load_cell_signals(a.X_in, a.cell_mols, a.mol_map, a.pos, s, a.m, a.H, a.W, base + i_next, lane);
        xa = v.x;
        xb = v.y;
        if (v.z != 0.0f) wflags |= 1u << MS_FLAG_BAD_POSITION;
      }
      cp_async_wait_all();
      __syncwarp();
      if (ok) {
        run_lanes(a, L, sb, cell, lane, cnt, wflags);
#pragma unroll 1
        for (int k = lane; k < s; k += 32) a.X_out[cell * s + k] = lds_f(sb + 4u * (unsigned)(L.oX0 + k));
      } else if (lane == 0) {
        a.defer_list[atomicAdd(a.defer_count, 1)] = (int)cell;  // integrated by the tier-2 launch
      }
      __syncwarp();
      i = i_next;
    }
  }
  if (lane < MS_N_FLAGS && ((wflags >> lane) & 1u)) sflags[lane] = 1u;
  __syncthreads();
  if (threadIdx.x < MS_N_FLAGS && sflags[threadIdx.x]) a.flags[threadIdx.x] = 1u;
}

template <int VEC>
__global__ void __launch_bounds__(MAX_THREADS, MS_TUNE_MIN_CTAS) kinetics_fused_kernel(const __grid_constant__ KinArgs a,
                                                                    float *dense_scratch) {
  extern __shared__ __align__(16) unsigned smem[];
  __shared__ unsigned sflags[MS_N_FLAGS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warps_per_cta = blockDim.x >> 5;
  if (threadIdx.x < MS_N_FLAGS) sflags[threadIdx.x] = 0;
  __syncthreads();
  // warp regions start at an address aligned to the largest array stride (layout comment)
  const unsigned raw = (unsigned)__cvta_generic_to_shared(smem);
  unsigned aligned = (raw + (unsigned)a.smem_align - 1u) & ~((unsigned)a.smem_align - 1u);
  // Opaque copy: without it the compiler recomputes this address (a dozen uniform instructions)
  // at every use inside the loops instead of keeping it.
  asm volatile("mov.b32 %0, %0;" : "+r"(aligned));
  unsigned *wbase = smem + ((aligned - raw) >> 2) + (size_t)warp * a.words_per_warp;
  const unsigned wshared = aligned + (unsigned)warp * (unsigned)a.words_per_warp * 4u;
  const int p = a.p, s = a.s, s32 = pow2_at_least_32(s);
  const long long gw = (long long)blockIdx.x * warps_per_cta + warp;
  float *dscr = dense_scratch + gw * (size_t)(3 * round32(p));
  const long long n_items = a.tier == 2 ? (long long)*a.defer_count : a.c;
  WarpMem w;
  if (a.sparse_enabled) w = carve(wbase, wshared, p, s, a.e_ent, a.e_nnz, a.e_reg);
  float *X0 = (float *)wbase, *X1 = X0 + s32, *Fsig = X1 + s32;

  // X0 of a cell: gathered (world.py:619-621) or read; lane i holds signals i, i+32 (s <= 64),
  // so that the next cell's values can be fetched while the current cell is integrated
  auto cell_of = [&](long long item) -> long long {
    return a.tier == 2 ? (long long)a.defer_list[item] : item;
  };
  auto load_x = [&](long long cell, int i) -> float {
    if (a.X_in) return a.X_in[cell * s + i];
    if (i < a.m) return a.cell_mols[cell * a.m + i];
    int px = a.pos[2 * cell], py = a.pos[2 * cell + 1];
    if ((unsigned)px >= (unsigned)a.H || (unsigned)py >= (unsigned)a.W) {
      px = py = 0;  // never read outside the map; reported through MS_FLAG_BAD_POSITION
      sflags[MS_FLAG_BAD_POSITION] = 1u;
    }
    return a.mol_map[((long long)(i - a.m) * a.H + px) * a.W + py];
  };
  if (a.batch_enabled) {
    // ---- batch mode: 8 cells per CTA at a time, scanned per warp, integrated together
    __shared__ BatchShared bs;
    __shared__ long long batch_cells[BATCH];
    const int tid = threadIdx.x;
    if (tid == 0) bs.next_base = (long long)atomicAdd(a.work_counter, BATCH);
    __syncthreads();
    long long base = bs.next_base;
    float xa = 0.0f, xb = 0.0f;
    if (base + warp < n_items) {
      if (lane < s) xa = load_x(base + warp, lane);
      if (lane + 32 < s) xb = load_x(base + warp, lane + 32);
    }
    __syncthreads();  // everyone has read next_base
    const unsigned region_bytes = a.region_bytes;
    const WarpMem w0 = carve(smem + ((aligned - raw) >> 2), aligned, p, s, a.e_ent, a.e_nnz, a.e_reg);
    while (base < n_items) {
      // The previous batch ended with a CTA barrier (end of run_batch), and everything written
      // here was last read before it.
      if (tid == 0) bs.next_base = (long long)atomicAdd(a.work_counter, BATCH);
      if (tid < 17) bs.cnt_p[tid] = bs.cur_p[tid] = 0;
      if (tid < 33) bs.cnt_c[tid] = bs.cur_c[tid] = 0;
      if (tid < BATCH) bs.cell_state[0][tid] = bs.cell_state[1][tid] = bs.cell_poison[tid] = 0;
      const long long cell = base + warp;
      const bool present = cell < n_items;
      int n_act = 0;
      bool ok = false;
      if (present) {
        if (lane < s) X0[lane] = xa;
        if (lane + 32 < s) X0[lane + 32] = xb;
        const long long prm = param_row(a, cell);
        const long long prow = prm * (long long)p;
        const float *gv = a.adj_vmax ? a.adj_vmax : a.Vmax;
        auto fetch_scalars = [&]() {  // per-protein scalars: asynchronous copies, awaited in build_lists
          for (int j = lane; j < p; j += 32) {
            cp_async4(w.aKe() + 4 * j, a.Ke + prow + j);
            cp_async4(w.aKmf() + 4 * j, a.Kmf + prow + j);
            cp_async4(w.aKmb() + 4 * j, a.Kmb + prow + j);
            cp_async4(w.aVmax() + 4 * j, gv + prow + j);
          }
        };
        // list cache: a valid record replaces the scan of the dense tile (the record holds exactly
        // what build_lists left in the region for these parameters, so the results are identical)
        unsigned *rec = a.cache ? a.cache + prm * (long long)a.rec_words : nullptr;
        const unsigned hdr = rec ? __ldg(rec) : 0u;
        const unsigned status = hdr >> 24;
        if (status == REC_VALID) {
          const uint4 *src = reinterpret_cast<const uint4 *>(rec + REC_HEADER_WORDS);
#pragma unroll
          for (int g = 0; g < 3; g++) {
            const unsigned dst = w.sb + 4u * (unsigned)a.rec_seg[2 * g];
            const int n16 = a.rec_seg[2 * g + 1] >> 2;
            for (int e = lane; e < n16; e += 32) cp_async16(dst + 16u * (unsigned)e, src + e, 0u);
            src += n16;
          }
          cp_async_wait_all();  // the pooling below reads act, rowptr and colptr
          __syncwarp();
          n_act = (int)(hdr & 0xffffu);
          ok = true;
        } else if (status != REC_OVERFLOW) {
          int nq;
          if (VEC == 4 && a.scan_staged) {
            if (a.scan_staged == 1) fetch_scalars();
            __syncwarp();
            nq = a.scan_staged == 2
                     ? scan_tile_staged<2>(a, w.aEnt(), w.aReg(), w.aX1(), w.aV(), w.aAct(), w.aKe(), w.aKmb(),
                                           w.aNvs(), a.e_nnz, prow * s, lane)
                     : scan_tile_staged<1>(a, w.aEnt(), w.aReg(), w.aX1(), w.aV(), w.aAct(), w.aKe(), w.aKmb(),
                                           w.aNvs(), a.e_nnz, prow * s, lane);
            if (a.scan_staged == 2) fetch_scalars();  // their arrays were staging slots until now
          } else {
            fetch_scalars();
            __syncwarp();
            nq = scan_tile<VEC>(a, w.queue(), a.e_nnz, prow * s, lane);
          }
          ok = nq >= 0 && build_lists<false>(a, w, prm, lane, nq, n_act);
          if (rec) {  // keep the lists for the next steps
            if (ok) {
              cp_async_wait_all();  // scalars and Kmr have landed
              __syncwarp();
              uint4 *dst = reinterpret_cast<uint4 *>(rec + REC_HEADER_WORDS);
#pragma unroll
              for (int g = 0; g < 3; g++) {
                const uint4 *seg = reinterpret_cast<const uint4 *>(w.b + a.rec_seg[2 * g]);
                const int n16 = a.rec_seg[2 * g + 1] >> 2;
                for (int e = lane; e < n16; e += 32) dst[e] = seg[e];
                dst += n16;
              }
            }
            if (lane == 0) rec[0] = (ok ? (REC_VALID << 24) | (unsigned)n_act : REC_OVERFLOW << 24);
          }
        }
        if (!ok) {
          cp_async_wait_all();
          if (a.defer_enabled) {
            if (lane == 0) a.defer_list[atomicAdd(a.defer_count, 1)] = (int)cell;
          } else {
            __syncwarp();
            run_dense(a, X0, X1, Fsig, dscr, dscr + round32(p), dscr + 2 * round32(p), cell, lane, sflags);
            for (int i = lane; i < s; i += 32) a.X_out[cell * s + i] = X0[i];
          }
        } else {
          for (int j = lane; j < p; j += 32) {  // run_batch's entry state
            sts_f(w.aV() + 4 * j, 0.0f);
            sts_f(w.aF() + 4 * j, 1.0f);
            sts_u(w.aFk() + 4 * j, KEY_ONE);
          }
        }
      }
      if (lane == 0) batch_cells[warp] = ok ? cell : -1;
      // pool the proteins and signals of the 8 cells, longest lists first (counting sort);
      // lane t holds protein act[t], act[t + 32] and signals t, t + 32
      unsigned aw_mine[2] = {0, 0}, bkt_p[2] = {0, 0}, bkt_c[2] = {0, 0};
      bool has_p[2] = {false, false}, has_c[2] = {false, false};
      if (ok) {
#pragma unroll
        for (int r = 0; r < 2; r++) {
          const int t = lane + 32 * r;
          if (t < n_act) {
            aw_mine[r] = w.act()[t];
            const unsigned j = (aw_mine[r] & 0x3fcu) >> 2;
            bkt_p[r] = 16u - min((w.rowptr()[j + 1] & 0xffffu) - (w.rowptr()[j] & 0xffffu), 16u);
            has_p[r] = true;
          }
          if (t < s) {
            bkt_c[r] = 32u - min(w.colptr()[t + 1] - w.colptr()[t], 32u);
            has_c[r] = true;
          }
        }
      }
      __syncthreads();  // counters are zero, next_base is visible
      const long long next_base = bs.next_base;
      if (next_base + warp < n_items) {  // next cell's signals: in flight during this batch
        if (lane < s) xa = load_x(next_base + warp, lane);
        if (lane + 32 < s) xb = load_x(next_base + warp, lane + 32);
      }
#pragma unroll
      for (int r = 0; r < 2; r++) {
        if (has_p[r]) atomicAdd(&bs.cnt_p[bkt_p[r]], 1u);
        if (has_c[r]) atomicAdd(&bs.cnt_c[bkt_c[r]], 1u);
      }
      __syncthreads();
      // every warp scans the two histograms for itself: lane b holds the first position of bucket b
      unsigned cp_ = lane < 17 ? bs.cnt_p[lane] : 0u, cc_ = bs.cnt_c[lane];
      unsigned ip = cp_, ic = cc_;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned yp = __shfl_up_sync(FULL, ip, d), yc = __shfl_up_sync(FULL, ic, d);
        if (lane >= d) {
          ip += yp;
          ic += yc;
        }
      }
      const int n_pitems = (int)__shfl_sync(FULL, ip, 31);
      const unsigned first32 = __shfl_sync(FULL, ic, 31);  // bucket 32 (empty lists) starts here
      const int n_citems = (int)(first32 + bs.cnt_c[32]);
#pragma unroll
      for (int r = 0; r < 2; r++) {
        const unsigned off_p = __shfl_sync(FULL, ip - cp_, (int)bkt_p[r]);
        unsigned off_c = __shfl_sync(FULL, ic - cc_, (int)(bkt_c[r] & 31u));
        if (bkt_c[r] == 32u) off_c = first32;
        if (has_p[r])
          bs.pitem[item_index(off_p + atomicAdd(&bs.cur_p[bkt_p[r]], 1u))] =
              (unsigned short)(((unsigned)warp << 12) | (aw_mine[r] & 0xfffu));
        if (has_c[r])
          bs.citem[item_index(off_c + atomicAdd(&bs.cur_c[bkt_c[r]], 1u))] =
              (unsigned short)(((unsigned)warp << 8) | (unsigned)(lane + 32 * r));
      }
      if (ok) {
        // The asynchronous copies of build_lists (scalars, Kmr) had the pooling above to land.
        cp_async_wait_all();
        __syncwarp();
        regulation_pass(w.sb, w.aRow() - w.sb, w.aReg() - w.sb, w.aRegk() - w.sb, w.aX1() - w.sb, p, lane);
      }
      __syncthreads();
      run_batch(a, w0, region_bytes, bs, batch_cells, n_pitems, n_citems, sflags);
      if (ok)
        for (int i = lane; i < s; i += 32) a.X_out[cell * s + i] = X0[i];
      base = next_base;
    }
    __syncthreads();
    if (threadIdx.x < MS_N_FLAGS && sflags[threadIdx.x]) a.flags[threadIdx.x] = 1u;
    return;
  }

  // Cells are handed out dynamically (one atomic per cell): with a static stride the slowest of
  // the 4736 warps finishes several per cent after the average one.  A warp always holds its
  // current item and the next one, so that the next cell's X0 can be fetched early.
  auto grab = [&]() -> long long {
    int v = 0;
    if (lane == 0) v = atomicAdd(a.work_counter, 1);
    return (long long)__shfl_sync(FULL, v, 0);
  };
  const bool prefetch = s <= 64;
  float xa = 0.0f, xb = 0.0f;
  long long item = grab();
  if (prefetch && item < n_items) {
    const long long c0 = cell_of(item);
    if (lane < s) xa = load_x(c0, lane);
    if (lane + 32 < s) xb = load_x(c0, lane + 32);
  }

  while (item < n_items) {
    const long long next_item = grab();
    const long long cell = cell_of(item);
    if (prefetch) {
      if (lane < s) X0[lane] = xa;
      if (lane + 32 < s) X0[lane + 32] = xb;
    } else {
      for (int i = lane; i < s; i += 32) X0[i] = load_x(cell, i);
    }
    const long long prm = param_row(a, cell);
    if (a.sparse_enabled) {  // per-protein scalars: asynchronous copies that land during the scan
      const long long prow = prm * (long long)p;
      const float *gv = a.adj_vmax ? a.adj_vmax : a.Vmax;
      for (int j = lane; j < p; j += 32) {
        cp_async4(w.aKe() + 4 * j, a.Ke + prow + j);
        cp_async4(w.aKmf() + 4 * j, a.Kmf + prow + j);
        cp_async4(w.aKmb() + 4 * j, a.Kmb + prow + j);
        cp_async4(w.aVmax() + 4 * j, gv + prow + j);
      }
    }
    __syncwarp();
    bool sparse = false;
    int n_act = 0;
    if (a.sparse_enabled) {
      const int nq = scan_tile<VEC>(a, w.queue(), a.e_nnz, prm * (long long)p * s, lane);
      sparse = nq >= 0 && build_lists<true>(a, w, prm, lane, nq, n_act);
      if (!sparse) cp_async_wait_all();
    }
    if (prefetch && next_item < n_items) {  // next cell's signals: in flight during this cell's parts
      const long long cn = cell_of(next_item);
      if (lane < s) xa = load_x(cn, lane);
      if (lane + 32 < s) xb = load_x(cn, lane + 32);
    }
    if (sparse) {
      run_sparse(a, w, cell, lane, n_act, sflags);
    } else if (a.tier == 1 && a.defer_enabled) {
      if (lane == 0) a.defer_list[atomicAdd(a.defer_count, 1)] = (int)cell;
      item = next_item;
      continue;  // integrated by the tier-2 launch
    } else {
      __syncwarp();
      run_dense(a, X0, X1, Fsig, dscr, dscr + round32(p), dscr + 2 * round32(p), cell, lane, sflags);
    }
    for (int i = lane; i < s; i += 32) a.X_out[cell * s + i] = X0[i];
    __syncwarp();
    item = next_item;
  }
  __syncthreads();
  if (threadIdx.x < MS_N_FLAGS && sflags[threadIdx.x]) a.flags[threadIdx.x] = 1u;
}

// ---------------------------------------------------------------------------
// staged kernels (dense tensors, one thread per (cell, protein) or (cell, signal))
// ---------------------------------------------------------------------------
__global__ void multiply_signals_kernel(const float *X, const int *K, long long c, int p, int s,
                                        float *xx, uint8_t *involved) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= c * p) return;
  bool inv;
  xx[r] = dense_product(X + (r / p) * s, K + r * s, s, inv);
  involved[r] = inv ? 1 : 0;
}

__global__ void quotient_kernel(const float *X, const int *Nf, const int *Nb, long long c, int p,
                                int s, float *Q) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= c * p) return;
  bool invb, invf;
  float xb = dense_product(X + (r / p) * s, Nb + r * s, s, invb);
  if (!invb) xb = 0.0f;
  float xf = dense_product(X + (r / p) * s, Nf + r * s, s, invf);
  if (!invf) xf = 0.0f;
  Q[r] = quotient_from(xb, xf);
}

__global__ void velocities_kernel(KinArgs a, const float *X, const float *vmax, float *V) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.c * a.p) return;
  const int s = a.s;
  const float *x = X + (r / a.p) * s;
  bool inv;
  float kf = dense_product(x, a.Nf + r * s, s, inv) / a.Kmf[r];
  if (!inv) kf = 0.0f;
  if (isinf(kf)) kf = V_MAX;
  float kb = dense_product(x, a.Nb + r * s, s, inv) / a.Kmb[r];
  if (!inv) kb = 0.0f;
  if (isinf(kb)) kb = V_MAX;
  float a_reg = 1.0f;
  for (int i = 0; i < s; i++) {
    const int av = a.A[r * s + i];
    if (av != 0) a_reg = a_reg * reg_factor(x[i], av, a.Kmr[r * s + i]);
  }
  V[r] = velocity_from(kf, kb, a_reg, vmax[r]);
}

// F_i = min(X_i / sum_j max(-NV_ji, 0), 1), one thread per (cell, signal); Fsig (c,s) scratch
__global__ void neg_adjust_fsig_kernel(const float *NV, const float *X, long long c, int p, int s,
                                       float *Fsig) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= c * s) return;
  const long long cell = r / s;
  const int i = (int)(r % s);
  float d = 0.0f;
  for (int j = 0; j < p; j++) {
    float neg = -NV[(cell * p + j) * s + i];
    neg = neg < 0.0f ? 0.0f : neg;
    d = d + neg;
  }
  float f = X[r] / d;
  if (f > 1.0f) f = 1.0f;
  Fsig[r] = f;
}

__global__ void neg_adjust_apply_kernel(const float *NV, const float *Fsig, long long c, int p,
                                        int s, float *NV_adj) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= c * p) return;
  const long long cell = r / p;
  float fm = 1.0f;
  for (int i = 0; i < s; i++)
    if (NV[r * s + i] < 0.0f) fm = nanmin(fm, Fsig[cell * s + i]);
  for (int i = 0; i < s; i++) NV_adj[r * s + i] = NV[r * s + i] * fm;
}

// one equilibrium iteration, decision half: one thread per (cell, protein)
__global__ void eq_decide_kernel(const float *X1, const float *V, const int *Nf, const int *Nb,
                                 const float *Ke, long long c, int p, int s, float *F, float inc,
                                 unsigned *flag_imp, unsigned *flag_any) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= c * p) return;
  const float *x = X1 + (r / p) * s;
  bool invb, invf;
  float xb = dense_product(x, Nb + r * s, s, invb);
  if (!invb) xb = 0.0f;
  float xf = dense_product(x, Nf + r * s, s, invf);
  if (!invf) xf = 0.0f;
  const float Fr = F[r];
  const unsigned dec = eq_decide(V[r], quotient_from(xb, xf), Ke[r], Fr);
  if (dec & 3u) {
    F[r] = eq_update(Fr, dec, inc);
    *flag_any = 1u;
    if (dec & 4u) *flag_imp = 1u;
  }
}

// X1_i = max(X0_i + sum_j NV_ji * F_j, 0), one thread per (cell, signal)
__global__ void eq_apply_kernel(const float *X0, const float *NV, const float *F, long long c, int p,
                                int s, float *X1) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= c * s) return;
  const long long cell = r / s;
  const int i = (int)(r % s);
  float acc = 0.0f;
  for (int j = 0; j < p; j++) acc = acc + NV[(cell * p + j) * s + i] * F[cell * p + j];
  X1[r] = clamp0(X0[r] + acc);
}

__global__ void fill_kernel(float *x, long long n, float v) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) x[r] = v;
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
struct LaunchPlan {
  int e_ent, e_nnz, e_reg, words_per_warp, sparse_enabled, grid, warps_per_cta, align;
  size_t smem_bytes;
};

// multiprocessor count of the current device (cached per device ordinal)
static int sm_count() {
  static int cached[64] = {0};
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev >= 0 && dev < 64 && cached[dev] > 0) return cached[dev];
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  if (dev >= 0 && dev < 64) cached[dev] = n;
  return n;
}

// Diagnostic switches, read from the environment ONCE, at the first launch of the
// process; ms_kinetics_reload_switches() re-reads them (tests flip them between launches).
constexpr int MAX_PHASE_BATCH = 1024, MAX_PHASED_CTAS = 192;
struct Switches {
  bool no_lanes, no_batch, no_staged_scan, staged_depth1, phased;
  int max_ctas_per_sm;  // 0 = no limit
  int phase_batch;      // cells per round and SM of the phased lane kernel
  int phase_ahead;      // how many cells ahead of the one a warp takes its tiles are prefetched into L2 (0: never)
};
static Switches read_switches() {
  Switches sw;
  sw.no_lanes = getenv("MS_KIN_NO_LANES") != nullptr;
  sw.no_batch = getenv("MS_KIN_NO_BATCH") != nullptr;
  sw.no_staged_scan = getenv("MS_KIN_NO_STAGED_SCAN") != nullptr;
  sw.staged_depth1 = getenv("MS_KIN_STAGED_DEPTH1") != nullptr;
  const char *env = getenv("MS_KIN_MAX_CTAS_PER_SM");
  sw.max_ctas_per_sm = env ? atoi(env) : 0;
  sw.phased = getenv("MS_KIN_PHASED") != nullptr;
  env = getenv("MS_KIN_PHASE_BATCH");
  sw.phase_batch = env ? atoi(env) : 512;
  if (sw.phase_batch < 32) sw.phase_batch = 32;
  if (sw.phase_batch > MAX_PHASE_BATCH) sw.phase_batch = MAX_PHASE_BATCH;
  env = getenv("MS_KIN_PHASE_AHEAD");
  sw.phase_ahead = env ? atoi(env) : 0;  // measured: any distance > 0 is slower
  if (sw.phase_ahead < 0) sw.phase_ahead = 0;
  if (sw.phase_ahead > 32) sw.phase_ahead = 32;
  return sw;
}
static Switches &switches() {
  static Switches sw = read_switches();
  return sw;
}

constexpr int MAX_GRID = 256 * 4;  // launches are clamped to this many CTAs (>= 4 per SM on any B200 part)
constexpr size_t SMEM_PER_SM = 222 * 1024;   // usable per SM (227 KB max per CTA on sm_100a)
constexpr size_t SMEM_PER_CTA_MAX = 200 * 1024;
constexpr size_t SMEM_PER_SM_HW = 227 * 1024;  // sm_100a: 228 KB per SM, 1 KB reserved per CTA

static bool sparse_possible(int64_t p, int64_t s) { return p >= 1 && p <= 256 && s >= 1 && s <= MAX_S_SPARSE; }

static int align_bytes(int64_t p, int64_t s) {
  const int S2 = pow2_at_least_32((int)s), P2 = pow2_at_least_32((int)p);
  return 4 * (S2 > P2 ? S2 : P2);
}

static void finish_plan(LaunchPlan &pl) {
  size_t per_sm = (SMEM_PER_SM_HW + 1024) / (pl.smem_bytes + 1024);
  if (per_sm > 4) per_sm = 4;
  if (per_sm < 1) per_sm = 1;
  // tuning knob: fewer resident CTAs per SM (the shared memory request is padded accordingly)
  if (switches().max_ctas_per_sm >= 1 && (size_t)switches().max_ctas_per_sm < per_sm) {
    per_sm = (size_t)switches().max_ctas_per_sm;
    const size_t pad = (SMEM_PER_SM_HW + 1024) / (per_sm + 1) + 1024;  // too big for per_sm + 1 CTAs
    if (pl.smem_bytes < pad && pad <= SMEM_PER_SM_HW) pl.smem_bytes = pad;
  }
  pl.grid = sm_count() * (int)per_sm;
  if (pl.grid > MAX_GRID) pl.grid = MAX_GRID;  // dense_scratch_bytes() is sized for this
}

// tier 1: lists for a typical proteome (2-6 non-zeros per protein row)
static LaunchPlan plan_tier1(int64_t p, int64_t s) {
  LaunchPlan pl = {};
  pl.sparse_enabled = sparse_possible(p, s) ? 1 : 0;
  const int full = (int)(p * s);
  // e_nnz also bounds the element queue (all non-zero positions of the tile)
  int e_ent = (int)((9 * p + 1) / 2), e_nnz = (int)((7 * p + 1) / 2), e_reg = (int)p;
  if (e_ent < 64) e_ent = 64;
  if (e_nnz < 64) e_nnz = 64;
  if (e_reg < 32) e_reg = 32;
  pl.e_ent = round32(e_ent < 2 * full ? e_ent : 2 * full);
  pl.e_nnz = round32(e_nnz < full ? e_nnz : full);
  pl.e_reg = round32(e_reg < full ? e_reg : full);
  pl.warps_per_cta = WARPS_PER_CTA;
  pl.align = align_bytes(p, s);
  pl.words_per_warp = make_layout((int)p, (int)s, pl.e_ent, pl.e_nnz, pl.e_reg).words;
  pl.smem_bytes = (size_t)pl.words_per_warp * 4 * pl.warps_per_cta + pl.align;
  if (!pl.sparse_enabled || pl.smem_bytes > SMEM_PER_CTA_MAX) {
    pl.sparse_enabled = 0;
    pl.e_ent = pl.e_nnz = pl.e_reg = 0;
    pl.align = align_bytes(1, s);  // the dense path keeps nothing of size p in shared memory
    pl.words_per_warp = 4 * pow2_at_least_32((int)s);
    pl.smem_bytes = (size_t)pl.words_per_warp * 4 * pl.warps_per_cta + pl.align;
  }
  finish_plan(pl);
  return pl;
}

static bool lanes_possible(int64_t p, int64_t s) {
  return p >= 1 && s >= 1 && p <= LANE_MAX_PS && s <= LANE_MAX_PS && !switches().no_lanes;
}

// tier 1 of the lane path (kinetics_lanes.cuh): one warp per cell, 8 independent warps per CTA
static LaunchPlan plan_lanes(int64_t p, int64_t s) {
  LaunchPlan pl = {};
  const LaneLayout L = make_lane_layout((int)p, (int)s);
  pl.sparse_enabled = 1;
  pl.e_nnz = L.q_cap;
  pl.warps_per_cta = WARPS_PER_CTA;
  pl.align = 128;
  pl.words_per_warp = L.words;
  pl.smem_bytes = (size_t)pl.words_per_warp * 4 * pl.warps_per_cta + pl.align;
  finish_plan(pl);
  return pl;
}

// tier 2: lists that hold a completely dense tile; as many warps per CTA as fit
static bool plan_tier2(int64_t p, int64_t s, LaunchPlan &pl) {
  pl = {};
  if (!sparse_possible(p, s)) return false;
  pl.sparse_enabled = 1;
  pl.e_nnz = pl.e_reg = round32((int)(p * s));
  pl.e_ent = 2 * pl.e_nnz;
  if (pl.e_ent > 0xffff) return false;  // 16-bit list offsets
  pl.align = align_bytes(p, s);
  pl.words_per_warp = make_layout((int)p, (int)s, pl.e_ent, pl.e_nnz, pl.e_reg).words;
  const size_t per_warp = (size_t)pl.words_per_warp * 4;
  if (per_warp + pl.align > SMEM_PER_CTA_MAX) return false;
  int w = (int)((SMEM_PER_SM / 2) / per_warp);
  if (w < 1) w = 1;
  if (w > WARPS_PER_CTA) w = WARPS_PER_CTA;
  pl.warps_per_cta = w;
  pl.smem_bytes = per_warp * w + pl.align;
  finish_plan(pl);
  return true;
}

static size_t align256k(size_t x) { return (x + 255) & ~(size_t)255; }

// scratch layout: [dense arrays: max_warps x 3 x round32(p) floats][defer counter][defer list: c ints]
static size_t dense_scratch_bytes(int64_t p) {
  return align256k((size_t)MAX_GRID * WARPS_PER_CTA * 3 * (size_t)round32((int)p) * sizeof(float));
}

}  // namespace ms

using namespace ms;

extern "C" void ms_kinetics_reload_switches(void) { switches() = read_switches(); }

// record slots of the phased lane kernel: one round per CTA
static size_t phased_scratch_bytes(int64_t p, int64_t s) {
  if (!lanes_possible(p, s) || !switches().phased) return 0;
  const size_t rec_words = REC_HEADER_WORDS + (size_t)make_lane_layout((int)p, (int)s).static_words;
  return align256k((size_t)MAX_PHASED_CTAS * (size_t)switches().phase_batch * rec_words * sizeof(unsigned));
}

extern "C" size_t ms_kinetics_scratch_bytes(int64_t c, int64_t p, int64_t s) {
  if (c < 0) c = 0;
  return dense_scratch_bytes(p) + 256 + align256k((size_t)c * sizeof(int)) + phased_scratch_bytes(p, s);
}

template <typename K>
static int launch_one(K kern, const LaunchPlan &pl, KinArgs &a, float *dense, cudaStream_t st) {
  MS_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
  a.e_ent = pl.e_ent; a.e_nnz = pl.e_nnz; a.e_reg = pl.e_reg; a.words_per_warp = pl.words_per_warp;
  a.region_bytes = (unsigned)pl.words_per_warp * 4u;
  a.smem_align = pl.align;
  a.batch_enabled = (a.tier == 1 && pl.sparse_enabled && a.p <= BATCH_MAX_PS && a.s <= BATCH_MAX_PS &&
                     pl.warps_per_cta == BATCH && !switches().no_batch) ? 1 : 0;
  {  // staging slots of scan_tile_staged: 1024 B in ent, 512 B in reg + regk, 512 B from X1 to colptr's end
    const Layout L = make_layout(a.p, a.s, pl.e_ent, pl.e_nnz, pl.e_reg);
    a.scan_staged = (a.batch_enabled && pl.e_ent >= 256 && 2 * pl.e_reg >= 128 && 3 * L.S2 + 32 >= 128 &&
                     !switches().no_staged_scan) ? 1 : 0;
    // second slot set (two steps in flight): 3*P2, 2*P2 + 32, 2*P2, 2*P2 words of 128
    if (a.scan_staged && L.P2 >= 64 && !switches().staged_depth1) a.scan_staged = 2;
  }
  a.sparse_enabled = pl.sparse_enabled;
  kern<<<pl.grid, pl.warps_per_cta * 32, pl.smem_bytes, st>>>(a, dense);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

// widest vector load the tile geometry and the pointers allow
static int pick_vec(const KinArgs &a) {
  const uintptr_t bits = (uintptr_t)a.N | (uintptr_t)a.Nf | (uintptr_t)a.Nb | (uintptr_t)a.A;
  if (a.s % 4 == 0 && (bits & 15) == 0) return 4;
  if (a.s % 2 == 0 && (bits & 7) == 0) return 2;
  return 1;
}

template <typename K>
static int launch_lanes_one(K kern, const LaunchPlan &pl, KinArgs &a, cudaStream_t st) {
  MS_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
  a.words_per_warp = pl.words_per_warp;
  a.region_bytes = (unsigned)pl.words_per_warp * 4u;
  a.smem_align = pl.align;
  a.lanes = make_lane_layout(a.p, a.s);
  a.sparse_enabled = 1;
  kern<<<pl.grid, pl.warps_per_cta * 32, pl.smem_bytes, st>>>(a);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

// one CTA of 32 warps per SM; false if the shape's regions do not fit 32 times
static bool launch_lanes_phased(const LaunchPlan &pl, KinArgs &a, cudaStream_t st, int &rc) {
  const size_t smem = (size_t)pl.words_per_warp * 4 * (PHASED_THREADS / 32) + pl.align;
  if (smem > SMEM_PER_SM_HW) return false;
  auto kern = kinetics_lanes_phased_kernel<4>;
  rc = MS_OK;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  a.words_per_warp = pl.words_per_warp;
  a.region_bytes = (unsigned)pl.words_per_warp * 4u;
  a.smem_align = pl.align;
  a.lanes = make_lane_layout(a.p, a.s);
  a.sparse_enabled = 1;
  a.phase_batch = switches().phase_batch;
  a.phase_ahead = switches().phase_ahead;
  int grid = sm_count();
  if (grid > MAX_PHASED_CTAS) grid = MAX_PHASED_CTAS;
  const long long rounds = ceil_div(a.c, a.phase_batch);
  if (grid > rounds) grid = (int)rounds;
  kern<<<grid, PHASED_THREADS, smem, st>>>(a);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    ms_set_cuda_error_(e);
    rc = MS_ERR_CUDA;
  }
  return true;
}

static int launch_lanes(const LaunchPlan &pl, KinArgs &a, cudaStream_t st) {
  int vec = pick_vec(a);
  if (vec == 4 && switches().no_staged_scan) vec = 2;  // register scan
  if (vec == 4 && switches().phased && !a.cache && a.phase_recs) {
    int rc = MS_OK;
    if (launch_lanes_phased(pl, a, st, rc)) return rc;
  }
  if (vec == 4) return launch_lanes_one(kinetics_lanes_kernel<4>, pl, a, st);
  if (vec == 2) return launch_lanes_one(kinetics_lanes_kernel<2>, pl, a, st);
  return launch_lanes_one(kinetics_lanes_kernel<1>, pl, a, st);
}

static int launch_plan(const LaunchPlan &pl, KinArgs &a, float *dense, cudaStream_t st) {
  const int vec = pick_vec(a);
  if (vec == 4) return launch_one(kinetics_fused_kernel<4>, pl, a, dense, st);
  if (vec == 2) return launch_one(kinetics_fused_kernel<2>, pl, a, dense, st);
  return launch_one(kinetics_fused_kernel<1>, pl, a, dense, st);
}

// words of a list-cache record for tier-1 lists of this shape (0: the shape has no batch mode)
static int64_t record_words(int64_t p, int64_t s, int *seg) {
  if (p < 1 || s < 1 || p > BATCH_MAX_PS || s > BATCH_MAX_PS) return 0;
  if (lanes_possible(p, s)) return REC_HEADER_WORDS + make_lane_layout((int)p, (int)s).static_words;
  const LaunchPlan pl = plan_tier1(p, s);
  if (!pl.sparse_enabled || pl.warps_per_cta != BATCH) return 0;
  const Layout L = make_layout((int)p, (int)s, pl.e_ent, pl.e_nnz, pl.e_reg);
  int sg[6] = {3 * L.S2, L.S2 + 32,                          // colptr
               L.oP + 3 * L.P2, 6 * L.P2 + 32 + pl.e_ent,    // Ke Kmf Kmb Vmax act rowptr ent
               L.oCsc, pl.e_nnz + 2 * pl.e_reg};             // csc reg regk
  if (seg)
    for (int i = 0; i < 6; i++) seg[i] = sg[i];
  return REC_HEADER_WORDS + sg[1] + sg[3] + sg[5];
}

static int launch_fused(const ms_kinetics_params *P, KinArgs &a, void *scratch, size_t scratch_bytes,
                        cudaStream_t st, const ms_kinetics_cache *cache = nullptr) {
  if (!P || P->c < 0 || P->p < 0 || P->s < 0) return MS_ERR_INVALID_ARGUMENT;
  if (P->p > 65535 || P->s > 8192 || P->c > INT32_MAX) return MS_ERR_TOO_LARGE;
  if (cache) {
    if ((cache->row_map || cache->records) && (a.adj_vmax || a.V_out)) return MS_ERR_INVALID_ARGUMENT;
    a.row_map = cache->row_map;
    if (cache->records) {
      const int64_t rw = record_words(P->p, P->s, a.rec_seg);
      if (rw == 0 || cache->rec_words != rw || ((uintptr_t)cache->records & 15)) return MS_ERR_INVALID_ARGUMENT;
      a.cache = cache->records;
      a.rec_words = (int)rw;
    }
  }
  a.N = P->N; a.Nf = P->Nf; a.Nb = P->Nb; a.A = P->A;
  a.Kmr = P->Kmr; a.Ke = P->Ke; a.Kmf = P->Kmf; a.Kmb = P->Kmb; a.Vmax = P->Vmax;
  a.c = P->c; a.p = (int)P->p; a.s = (int)P->s;
  MS_CUDA_CHECK(cudaMemsetAsync(a.flags, 0, MS_N_FLAGS * sizeof(unsigned), st));
  if (P->c == 0 || P->s == 0) return MS_OK;
  if (!scratch || scratch_bytes < ms_kinetics_scratch_bytes(P->c, P->p, P->s)) return MS_ERR_SCRATCH_TOO_SMALL;
  float *dense = (float *)scratch;
  a.defer_count = (int *)((char *)scratch + dense_scratch_bytes(P->p));
  a.work_counter = a.defer_count + 1;  // [0] deferred count, [1] tier-1 work, [2] tier-2 work
  a.defer_list = a.defer_count + 64;
  a.phase_recs = phased_scratch_bytes(P->p, P->s)
                     ? (unsigned *)((char *)scratch + dense_scratch_bytes(P->p) + 256 + align256k((size_t)P->c * sizeof(int)))
                     : nullptr;

  // the lane kernel defers what it cannot pack, so it needs a tier 2 (true for every p, s <= 64)
  LaunchPlan p2;
  const bool lanes = lanes_possible(P->p, P->s) && plan_tier2(P->p, P->s, p2);
  LaunchPlan p1 = lanes ? plan_lanes(P->p, P->s) : plan_tier1(P->p, P->s);
  const bool two_tiers = lanes || (p1.sparse_enabled && plan_tier2(P->p, P->s, p2) && p2.e_nnz > p1.e_nnz);
  const int64_t ctas_needed = ceil_div(P->c, p1.warps_per_cta);
  if (p1.grid > ctas_needed) p1.grid = (int)ctas_needed;
  a.tier = 1;
  a.defer_enabled = two_tiers ? 1 : 0;
  MS_CUDA_CHECK(cudaMemsetAsync(a.defer_count, 0, 4 * sizeof(int), st));
  int rc = lanes ? launch_lanes(p1, a, st) : launch_plan(p1, a, dense, st);
  if (rc != MS_OK || !two_tiers) return rc;
  a.tier = 2;
  a.defer_enabled = 0;
  a.work_counter = a.defer_count + 2;
  const int64_t ctas2 = ceil_div(P->c, p2.warps_per_cta);
  if (p2.grid > ctas2) p2.grid = (int)ctas2;
  return launch_plan(p2, a, dense, st);
}

static int fill_eq(KinArgs &a, const int32_t *eq_iters_host) {
  for (int q = 0; q < 3; q++) {
    int v = eq_iters_host ? eq_iters_host[q] : 4;
    if (v < 0 || v > 4) return MS_ERR_INVALID_ARGUMENT;
    a.eq_iters[q] = v;
  }
  a.trims[0] = 0.7f; a.trims[1] = 0.2f; a.trims[2] = 0.1f;  // kinetics.py:746
  a.n_parts = 3;
  return MS_OK;
}

extern "C" int ms_integrate_signals(const ms_kinetics_params *params, const float *X_in, float *X_out,
                                    const int32_t *eq_iters_host, uint32_t *flags, void *scratch,
                                    size_t scratch_bytes, void *stream) {
  if (!params || !flags || (params->c > 0 && (!X_in || !X_out)) || X_in == X_out)
    return MS_ERR_INVALID_ARGUMENT;
  KinArgs a = {};
  int rc = fill_eq(a, eq_iters_host);
  if (rc) return rc;
  a.X_in = X_in; a.X_out = X_out; a.flags = flags;
  return launch_fused(params, a, scratch, scratch_bytes, (cudaStream_t)stream);
}

extern "C" int64_t ms_kinetics_cache_record_words(int64_t p, int64_t s) { return record_words(p, s, nullptr); }

__global__ void cache_invalidate_kernel(unsigned *records, long long rec_words, const int *rows, long long n) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) records[(rows ? (long long)rows[r] : r) * rec_words] = 0u;  // REC_STALE
}

extern "C" int ms_kinetics_cache_invalidate(uint32_t *records, int64_t rec_words, const int32_t *rows,
                                            int64_t n_rows, void *stream) {
  if (!records || rec_words < REC_HEADER_WORDS || n_rows < 0) return MS_ERR_INVALID_ARGUMENT;
  if (n_rows == 0) return MS_OK;
  cache_invalidate_kernel<<<(unsigned)((n_rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(records, rec_words,
                                                                                              rows, n_rows);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_integrate_signals_cached(const ms_kinetics_params *params, const ms_kinetics_cache *cache,
                                           const float *X_in, float *X_out, const int32_t *eq_iters_host,
                                           uint32_t *flags, void *scratch, size_t scratch_bytes, void *stream) {
  if (!params || !cache || !flags || (params->c > 0 && (!X_in || !X_out)) || X_in == X_out)
    return MS_ERR_INVALID_ARGUMENT;
  KinArgs a = {};
  int rc = fill_eq(a, eq_iters_host);
  if (rc) return rc;
  a.X_in = X_in; a.X_out = X_out; a.flags = flags;
  return launch_fused(params, a, scratch, scratch_bytes, (cudaStream_t)stream, cache);
}

extern "C" int ms_integrate_signals_part(const ms_kinetics_params *params, const float *adj_vmax,
                                         const float *X0, float *X1, int32_t max_eq_iters,
                                         uint32_t *flags, float *V_out, void *scratch,
                                         size_t scratch_bytes, void *stream) {
  if (!params || !flags || (params->c > 0 && (!X0 || !X1 || !adj_vmax)) || X0 == X1 ||
      max_eq_iters < 0 || max_eq_iters > 4)
    return MS_ERR_INVALID_ARGUMENT;
  KinArgs a = {};
  a.n_parts = 1;
  a.eq_iters[0] = max_eq_iters;
  a.trims[0] = 1.0f;
  a.adj_vmax = adj_vmax;
  a.V_out = V_out;
  a.X_in = X0; a.X_out = X1; a.flags = flags;
  return launch_fused(params, a, scratch, scratch_bytes, (cudaStream_t)stream);
}

static int enzymatic_activity_impl(const ms_kinetics_params *params, const ms_kinetics_cache *cache,
                                   const int32_t *cell_positions, const float *cell_molecules,
                                   const float *molecule_map, int64_t m, int64_t H, int64_t W, float *X_out,
                                   const int32_t *eq_iters_host, uint32_t *flags, void *scratch,
                                   size_t scratch_bytes, void *stream) {
  if (!params || !flags || params->s != 2 * m || H <= 0 || W <= 0 || H > INT32_MAX || W > INT32_MAX ||
      H * W > INT32_MAX ||
      (params->c > 0 && (!cell_positions || !cell_molecules || !molecule_map || !X_out)))
    return MS_ERR_INVALID_ARGUMENT;
  KinArgs a = {};
  int rc = fill_eq(a, eq_iters_host);
  if (rc) return rc;
  a.X_in = nullptr; a.cell_mols = cell_molecules; a.mol_map = molecule_map; a.pos = cell_positions;
  a.m = (int)m; a.H = (int)H; a.W = (int)W;
  a.X_out = X_out; a.flags = flags;
  return launch_fused(params, a, scratch, scratch_bytes, (cudaStream_t)stream, cache);
}

extern "C" int ms_enzymatic_activity(const ms_kinetics_params *params, const int32_t *cell_positions,
                                     const float *cell_molecules, const float *molecule_map,
                                     int64_t m, int64_t H, int64_t W, float *X_out,
                                     const int32_t *eq_iters_host, uint32_t *flags, void *scratch,
                                     size_t scratch_bytes, void *stream) {
  return enzymatic_activity_impl(params, nullptr, cell_positions, cell_molecules, molecule_map, m, H, W, X_out,
                                 eq_iters_host, flags, scratch, scratch_bytes, stream);
}

extern "C" int ms_enzymatic_activity_cached(const ms_kinetics_params *params, const ms_kinetics_cache *cache,
                                            const int32_t *cell_positions, const float *cell_molecules,
                                            const float *molecule_map, int64_t m, int64_t H, int64_t W,
                                            float *X_out, const int32_t *eq_iters_host, uint32_t *flags,
                                            void *scratch, size_t scratch_bytes, void *stream) {
  if (!cache) return MS_ERR_INVALID_ARGUMENT;
  return enzymatic_activity_impl(params, cache, cell_positions, cell_molecules, molecule_map, m, H, W, X_out,
                                 eq_iters_host, flags, scratch, scratch_bytes, stream);
}

static inline unsigned blocks_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

extern "C" int ms_multiply_signals(const float *X, const int32_t *K, int64_t c, int64_t p, int64_t s,
                                   float *xx, uint8_t *involved, void *stream) {
  if (c < 0 || p < 0 || s < 0) return MS_ERR_INVALID_ARGUMENT;
  if (c * p == 0) return MS_OK;
  multiply_signals_kernel<<<blocks_for(c * p, 128), 128, 0, (cudaStream_t)stream>>>(X, K, c, (int)p, (int)s, xx, involved);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_get_quotient(const float *X, const int32_t *Nf, const int32_t *Nb, int64_t c,
                               int64_t p, int64_t s, float *Q, void *stream) {
  if (c < 0 || p < 0 || s < 0) return MS_ERR_INVALID_ARGUMENT;
  if (c * p == 0) return MS_OK;
  quotient_kernel<<<blocks_for(c * p, 128), 128, 0, (cudaStream_t)stream>>>(X, Nf, Nb, c, (int)p, (int)s, Q);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_get_velocities(const ms_kinetics_params *P, const float *X, const float *adj_vmax,
                                 float *V, void *stream) {
  if (!P) return MS_ERR_INVALID_ARGUMENT;
  if (P->c * P->p == 0) return MS_OK;
  KinArgs a = {};
  a.Nf = P->Nf; a.Nb = P->Nb; a.A = P->A; a.Kmr = P->Kmr; a.Kmf = P->Kmf; a.Kmb = P->Kmb;
  a.c = P->c; a.p = (int)P->p; a.s = (int)P->s;
  velocities_kernel<<<blocks_for(P->c * P->p, 128), 128, 0, (cudaStream_t)stream>>>(a, X, adj_vmax, V);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_get_negative_adjusted_nv(const float *NV, const float *X, int64_t c, int64_t p,
                                           int64_t s, float *NV_adj, float *Fsig_scratch,
                                           void *stream) {
  if (c < 0 || p < 0 || s < 0) return MS_ERR_INVALID_ARGUMENT;
  if (c * p * s == 0) return MS_OK;
  if (!Fsig_scratch || NV == NV_adj) return MS_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  neg_adjust_fsig_kernel<<<blocks_for(c * s, 128), 128, 0, st>>>(NV, X, c, (int)p, (int)s, Fsig_scratch);
  MS_CUDA_CHECK(cudaGetLastError());
  neg_adjust_apply_kernel<<<blocks_for(c * p, 128), 128, 0, st>>>(NV, Fsig_scratch, c, (int)p, (int)s, NV_adj);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_get_equilibrium_adjusted_x(const float *X0, const float *X1_in, const float *NV,
                                             const float *V, const int32_t *Nf, const int32_t *Nb,
                                             const float *Ke, int64_t c, int64_t p, int64_t s,
                                             float *X1_out, float *F_scratch, int32_t max_iters,
                                             uint32_t *flags, void *stream) {
  if (c < 0 || p < 0 || s < 0 || max_iters < 0 || max_iters > 4 || !flags) return MS_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  MS_CUDA_CHECK(cudaMemsetAsync(flags, 0, MS_N_FLAGS * sizeof(unsigned), st));
  if (c * s == 0) return MS_OK;
  if (X1_out != X1_in)
    MS_CUDA_CHECK(cudaMemcpyAsync(X1_out, X1_in, sizeof(float) * (size_t)(c * s), cudaMemcpyDeviceToDevice, st));
  if (c * p == 0) return MS_OK;
  if (!F_scratch) return MS_ERR_INVALID_ARGUMENT;
  fill_kernel<<<blocks_for(c * p, 256), 256, 0, st>>>(F_scratch, c * p, 1.0f);
  for (int k = 0; k < max_iters; k++) {
    eq_decide_kernel<<<blocks_for(c * p, 128), 128, 0, st>>>(X1_out, V, Nf, Nb, Ke, c, (int)p, (int)s, F_scratch,
                                                             k == 0 ? 0.5f : (k == 1 ? 0.25f : (k == 2 ? 0.125f : 0.0625f)),
                                                             flags + k, flags + 12 + k);
    eq_apply_kernel<<<blocks_for(c * s, 128), 128, 0, st>>>(X0, NV, F_scratch, c, (int)p, (int)s, X1_out);
  }
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_add_protein_rates(const float *X0, const float *NV, int64_t c, int64_t p, int64_t s,
                                    float *X1, float *F_scratch, void *stream) {
  if (c < 0 || p < 0 || s < 0) return MS_ERR_INVALID_ARGUMENT;
  if (c * s == 0) return MS_OK;
  if (!X0 || !X1 || (c * p > 0 && (!NV || !F_scratch))) return MS_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  if (c * p > 0) fill_kernel<<<blocks_for(c * p, 256), 256, 0, st>>>(F_scratch, c * p, 1.0f);
  eq_apply_kernel<<<blocks_for(c * s, 128), 128, 0, st>>>(X0, NV, F_scratch, c, (int)p, (int)s, X1);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}
