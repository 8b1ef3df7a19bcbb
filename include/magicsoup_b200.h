/*
 * magicsoup_b200 -- C ABI of the B200 (sm_100a) implementation of magicsoup's
 * per-step world math.  One shared library: magicsoup_b200/libmagicsoup_b200.so.
 *
 * Conventions
 *   - Every pointer is a DEVICE pointer owned by the caller unless its name ends
 *     in `_host`.  The library allocates nothing and frees nothing; scratch
 *     memory is passed in (sizes from the ms_*_scratch_bytes queries).
 *   - Tensors are row-major and shaped exactly like the reference's torch tensors
 *     (reference = mRcSchwering/magic-soup 1.0.5, paths below are relative to its
 *     python/magicsoup/): X (c,s) f32; N,Nf,Nb,A (c,p,s) int32; Kmr (c,p,s) f32;
 *     Ke,Kmf,Kmb,Vmax (c,p) f32; molecule_map (m,H,W) f32; cell_molecules (c,m)
 *     f32; cell_positions (c,2) int32.  s = 2*m, intracellular signals first.
 *   - `stream` is a cudaStream_t (NULL = default stream).  Calls only enqueue
 *     work; nothing synchronises the device.
 *   - Return value: 0 = ok, negative = error (ms_error_string).  No C++ exceptions
 *     cross this boundary, NaN/Inf in the data never trap (the reference lets them
 *     propagate, docs/tutorials.md:408).
 *   - Arithmetic: float32, no FMA contraction, sums in ascending index order,
 *     integer powers by repeated multiplication -- bit-identical to
 *     oracle/ms_oracle.c (the "spec order").
 */
#ifndef MAGICSOUP_B200_H
#define MAGICSOUP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MS_OK 0
#define MS_ERR_INVALID_ARGUMENT (-1)
#define MS_ERR_CUDA (-2)
#define MS_ERR_SCRATCH_TOO_SMALL (-3)
#define MS_ERR_NO_DEVICE (-4)
#define MS_ERR_TOO_LARGE (-5)

/* Number of uint32 words in a flags buffer (see ms_integrate_signals). */
#define MS_N_FLAGS 32
/* flags[MS_FLAG_BAD_POSITION] != 0: some cell_positions entry lay outside the map (the gather read
 * pixel (0,0) instead; the reference raises IndexError, world.py:621) */
#define MS_FLAG_BAD_POSITION 24

/* The nine Kinetics parameter tensors, kinetics.py:399-409. */
typedef struct ms_kinetics_params {
  const int32_t *N, *Nf, *Nb, *A; /* (c,p,s) */
  const float *Kmr;               /* (c,p,s) */
  const float *Ke, *Kmf, *Kmb, *Vmax; /* (c,p) */
  int64_t c, p, s;
} ms_kinetics_params;

int ms_version(void);
const char *ms_error_string(int code);
/* message of the last CUDA error seen by this thread's failing call ("" if none) */
const char *ms_last_cuda_error(void);
int ms_device_count(void);

/* ------------------------------------------------------------------ kinetics */

/*
 * Kinetics.integrate_signals(X) -> X'          kinetics.py:725-751
 * Three trimmed parts (0.7, 0.2, 0.1), each: velocities, substrate trimming,
 * equilibrium trimming; fused per cell, every parameter read once.
 *
 * eq_iters_host[3]: how many equilibrium-adjustment iterations to APPLY in each
 *   part (0..4).  The reference stops a part's iterations as soon as NO cell in
 *   the whole batch has an adjustable protein with |V| > 0.1 (kinetics.py:846, a
 *   global torch.any).  Pass {4,4,4} first, then inspect `flags`:
 *     flags[q*4+k]      != 0  <=> some cell had such a protein at iteration k of part q
 *     flags[12+q*4+k]   != 0  <=> some cell had ANY adjustable protein there
 *   If flags[q*4+k]==0 while flags[12+q*4+k]!=0 for some k < eq_iters[q], the
 *   reference would have stopped part q at k: call again with eq_iters[q] = k
 *   (later parts' flags are only meaningful once earlier parts are settled).
 *   magicsoup_b200.kinetics.Kinetics.integrate_signals runs this protocol.
 * flags: MS_N_FLAGS uint32, zeroed by this call before the kernel runs.
 * scratch: ms_kinetics_scratch_bytes(c, p, s) bytes: the list of cells whose tiles do not
 *   fit the small in-shared-memory lists (they are integrated by a second launch of the
 *   same kernel with large lists) and the per-warp arrays of the dense path.
 * X_out must not alias X_in.
 */
int ms_integrate_signals(const ms_kinetics_params *params, const float *X_in, float *X_out,
                         const int32_t *eq_iters_host, uint32_t *flags, void *scratch,
                         size_t scratch_bytes, void *stream);

/*
 * Kinetics._integrate_signals_part(adj_vmax, X0) -> X1      kinetics.py:753-769
 * One part with the caller's already trimmed velocities adj_vmax (c,p).
 * max_eq_iters as eq_iters above (0 = no equilibrium adjustment, which is what the
 * reference's tests mock, tests/fast/test_kinetics.py:87-96).  flags: MS_N_FLAGS
 * words, entries [k] and [12+k] are written.  V_out (c,p) optional (NULL).
 */
int ms_integrate_signals_part(const ms_kinetics_params *params, const float *adj_vmax,
                              const float *X0, float *X1, int32_t max_eq_iters,
                              uint32_t *flags, float *V_out, void *scratch,
                              size_t scratch_bytes, void *stream);

size_t ms_kinetics_scratch_bytes(int64_t c, int64_t p, int64_t s);

/* The fused kernel's diagnostic switches (environment variables MS_KIN_NO_BATCH,
 * MS_KIN_NO_STAGED_SCAN, MS_KIN_STAGED_DEPTH1, MS_KIN_MAX_CTAS_PER_SM) are read
 * once, at the first launch of the process.  This re-reads them (no counterpart in the
 * reference; used by the tests that run every kernel path). */
void ms_kinetics_reload_switches(void);

/*
 * Parameter rows and the list cache -- no counterpart in the reference, which re-reads the dense
 * tensors with every aten op (kinetics.py:753-918) and re-allocates them on every cell-count
 * change (kinetics.py:667-703).
 *   row_map (optional): row_map[i] is the row of the nine parameter tensors (and of `records`)
 *     that holds logical cell i, so that a store can remove and append cells without moving
 *     42 KB per cell; X, cell_positions and cell_molecules stay indexed by the logical cell.
 *     params->c is the logical cell count; the tensors must have a row for every mapped value.
 *   records (optional): one record of rec_words = ms_kinetics_cache_record_words(p, s) uint32 per
 *     parameter ROW, 16-byte aligned, holding the sparse lists the kernel builds from that row's
 *     dense tile.  A row whose record is valid skips the scan of its 20*p*s bytes; results are
 *     bit-identical either way.  Word 0 of a record is its status: zero-fill a record
 *     (ms_kinetics_cache_invalidate, or any memset) whenever ANY of the row's parameters change.
 *     Only shapes with p, s <= 64 have records (ms_kinetics_cache_record_words returns 0 otherwise).
 */
typedef struct ms_kinetics_cache {
  uint32_t *records;
  int64_t rec_words;
  const int32_t *row_map;
} ms_kinetics_cache;

int64_t ms_kinetics_cache_record_words(int64_t p, int64_t s);
/* rows: n_rows parameter rows whose records become stale (NULL = rows 0 .. n_rows-1) */
int ms_kinetics_cache_invalidate(uint32_t *records, int64_t rec_words, const int32_t *rows,
                                 int64_t n_rows, void *stream);
/* ms_integrate_signals / ms_enzymatic_activity (below) with a row map and/or a list cache */
int ms_integrate_signals_cached(const ms_kinetics_params *params, const ms_kinetics_cache *cache,
                                const float *X_in, float *X_out, const int32_t *eq_iters_host,
                                uint32_t *flags, void *scratch, size_t scratch_bytes, void *stream);
int ms_enzymatic_activity_cached(const ms_kinetics_params *params, const ms_kinetics_cache *cache,
                                 const int32_t *cell_positions, const float *cell_molecules,
                                 const float *molecule_map, int64_t m, int64_t H, int64_t W, float *X_out,
                                 const int32_t *eq_iters_host, uint32_t *flags, void *scratch,
                                 size_t scratch_bytes, void *stream);

/*
 * Kinetics.set_cell_params(cell_idxs, proteomes)               kinetics.py:521-625
 * Domain tokens -> the nine parameter tensors, written into the given parameter rows.
 *   tables: the reference's token maps as device arrays: vmax_map.weights, km_map.weights
 *     (float, NaN at token 0), sign_map.signs, hill_map.numbers (int32), reaction_map.M,
 *     transport_map.M, effector_map.M (int32, (tokens, s), zero vector at token 0)
 *     (kinetics.py:16-289, 418-449);
 *   proteomes flattened (device, int32): rows[k] = parameter row of proteome k; its proteins are
 *     prot_offsets[k] .. prot_offsets[k+1]; protein q's domains are dom_offsets[q] ..
 *     dom_offsets[q+1]; doms[5*d ..] = (domain type, i0, i1, i2, i3) (constants.py:9);
 *   every protein slot 0 .. p-1 of the rows is written (slots beyond a proteome get the
 *     reference's padding: N = Nf = Nb = A = 0, Vmax = 0, Ke = 1, Kmf = Kmb = 1e-36, Kmr = 1).
 * Integer results are exact; float results follow the reference's float32 formulas (the energy
 * sum of kinetics.py:612 is accumulated in float64).  Tokens must index inside the tables.
 */
typedef struct ms_token_tables {
  const float *vmax_weights, *km_weights;
  const int32_t *signs, *hill_numbers;
  const int32_t *reaction_M, *transport_M, *effector_M;
} ms_token_tables;

typedef struct ms_kinetics_params_out {
  int32_t *N, *Nf, *Nb, *A; /* (rows,p,s) */
  float *Kmr;               /* (rows,p,s) */
  float *Ke, *Kmf, *Kmb, *Vmax; /* (rows,p) */
  int64_t p, s;
} ms_kinetics_params_out;

int ms_set_cell_params(const ms_token_tables *tables, const float *mol_energies, float abs_temp,
                       const int32_t *rows, const int32_t *prot_offsets, const int32_t *dom_offsets,
                       const int32_t *doms, int64_t n_cells, const ms_kinetics_params_out *out,
                       void *stream);

/* Kinetics._multiply_signals(X, N) -> (xx (c,p), involved (c,p) as uint8)  kinetics.py:894-918 */
int ms_multiply_signals(const float *X, const int32_t *K, int64_t c, int64_t p, int64_t s,
                        float *xx, uint8_t *involved, void *stream);

/* Kinetics._get_quotient(X) -> Q (c,p)                                      kinetics.py:881-892 */
int ms_get_quotient(const float *X, const int32_t *Nf, const int32_t *Nb, int64_t c, int64_t p,
                    int64_t s, float *Q, void *stream);

/* Kinetics._get_velocities(X, Vmax) -> V (c,p); Vmax is adj_vmax           kinetics.py:771-806 */
int ms_get_velocities(const ms_kinetics_params *params, const float *X, const float *adj_vmax,
                      float *V, void *stream);

/* Kinetics._get_negative_adjusted_nv(NV, X) -> NV_adj (c,p,s)              kinetics.py:861-879
 * Fsig_scratch: (c,s) floats of caller scratch.  NV_adj must not alias NV. */
int ms_get_negative_adjusted_nv(const float *NV, const float *X, int64_t c, int64_t p, int64_t s,
                                float *NV_adj, float *Fsig_scratch, void *stream);

/* X1 = max(X0 + sum over proteins of NV, 0)                                  kinetics.py:763-764
 * NV dense (c,p,s); F_scratch: (c,p) floats of caller scratch. */
int ms_add_protein_rates(const float *X0, const float *NV, int64_t c, int64_t p, int64_t s,
                         float *X1, float *F_scratch, void *stream);

/*
 * Kinetics._get_equilibrium_adjusted_x(X0, X1, NV, V) -> X1'               kinetics.py:808-859
 * Dense NV (c,p,s).  F (c,p) is caller scratch (float).  max_iters / flags as in
 * ms_integrate_signals_part.  X1_out may alias X1_in.
 */
int ms_get_equilibrium_adjusted_x(const float *X0, const float *X1_in, const float *NV,
                                  const float *V, const int32_t *Nf, const int32_t *Nb,
                                  const float *Ke, int64_t c, int64_t p, int64_t s, float *X1_out,
                                  float *F_scratch, int32_t max_iters, uint32_t *flags,
                                  void *stream);

/* --------------------------------------------------------------------- world */

/*
 * World.enzymatic_activity(), compute half                     world.py:610-622
 * Gathers X0 = [cell_molecules | molecule_map at the cell's pixel] inside the
 * kinetics kernel and writes the integrated signals to X_out (c,s).  Nothing is
 * modified in place so that the eq_iters protocol can re-run; commit the result
 * with ms_scatter_signals once `flags` are settled.
 */
int ms_enzymatic_activity(const ms_kinetics_params *params, const int32_t *cell_positions,
                          const float *cell_molecules, const float *molecule_map, int64_t m,
                          int64_t H, int64_t W, float *X_out, const int32_t *eq_iters_host,
                          uint32_t *flags, void *scratch, size_t scratch_bytes, void *stream);

/*
 * World.enzymatic_activity(), write-back half                  world.py:624-625
 * molecule_map[:, x, y] = X[:, m:], cell_molecules_out = X[:, :m].
 */
int ms_scatter_signals(const float *X, const int32_t *cell_positions, int64_t c, int64_t m,
                       int64_t H, int64_t W, float *cell_molecules_out, float *molecule_map,
                       void *stream);

/*
 * World.diffuse_molecules(), diffusion half                    world.py:638-649
 * Per molecule k: 3x3 circular stencil with weights a[k] (8 neighbours) and b[k]
 * (centre) from World._get_diffuse (world.py:948-972), then the mass correction
 * += (sum_before - sum_after) / (H*W) and clamp(0).  In place on molecule_map.
 * scratch: ms_diffuse_scratch_bytes(m,H,W) bytes (stencil output + plane sums).
 * Value range: the plane sums are exact 2^-32 fixed point, which holds finite |x| < 2^31; a NaN,
 * Inf or larger value turns its whole plane NaN (NaN / Inf: as the reference's float32 torch.sum
 * does, world.py:640-648).  Pointers need only float alignment; the
 * float4 paths are taken when the buffers are 16-byte aligned.
 */
int ms_diffuse_molecules(float *molecule_map, const float *a, const float *b, int64_t m, int64_t H,
                         int64_t W, void *scratch, size_t scratch_bytes, void *stream);

size_t ms_diffuse_scratch_bytes(int64_t m, int64_t H, int64_t W);

/*
 * The two passes of ms_diffuse_molecules for a ROW SLAB of a map that is split over
 * several GPUs (no counterpart in the single-device reference; same arithmetic as
 * world.py:638-649 on the whole map).
 * ms_diffuse_stencil: stencil of the slab (m,H,W) into `out` (m,H,W); the rows above
 *   row 0 and below row H-1 come from halo_top / halo_bottom (m,W) (the neighbouring
 *   slabs' edge rows; NULL = wrap around inside the slab); sums (5*m int64, zeroed by
 *   the call) receive the slab's exact plane sums before/after as 2^-32 fixed point,
 *   {hi_before, lo_before, hi_after, lo_after} per plane in the first 4*m words, and in word
 *   4*m + k the number of blocks that met a value the fixed point cannot hold in plane k (NaN,
 *   Inf, |x| >= 2^31): add all 5*m words over all slabs (integers, so the order does not
 *   matter) before the second pass.
 * ms_diffuse_correct: map_slab = clamp(out + (sum_before - sum_after)/n_pxls_total, 0); a plane
 *   whose flag word is non-zero becomes NaN, as with the reference's float32 torch.sum.
 */
int ms_diffuse_stencil(const float *map_slab, const float *halo_top, const float *halo_bottom,
                       const float *a, const float *b, int64_t m, int64_t H, int64_t W, float *out,
                       int64_t *sums, void *stream);
int ms_diffuse_correct(const float *out, float *map_slab, const int64_t *sums, int64_t m, int64_t H,
                       int64_t W, int64_t n_pxls_total, void *stream);

/*
 * Pixel-ordered write-back and permeation for densely populated maps -- same results as
 * ms_scatter_signals (world.py:624-625) and ms_permeate / ms_degrade_permeate (world.py:651-665),
 * with coalesced map accesses: one thread per pixel walks the m planes.
 * ms_build_pixel_index fills pixel_cell (H*W int32) with the index of the cell at each pixel, -1
 * where there is none (the reference's cell_map, world.py:194, with the index instead of a bool);
 * rebuild it whenever cell_positions or the cell order change.  decay may be NULL.  m <= 32
 * (MS_ERR_TOO_LARGE otherwise).
 */
int ms_build_pixel_index(const int32_t *cell_positions, int64_t c, int64_t H, int64_t W, int32_t *pixel_cell,
                         void *stream);
int ms_scatter_signals_by_pixel(const float *X, const int32_t *pixel_cell, int64_t m, int64_t H, int64_t W,
                                float *cell_molecules_out, float *molecule_map, void *stream);
int ms_permeate_by_pixel(const float *cell_molecules_in, float *cell_molecules_out, float *molecule_map,
                         const int32_t *pixel_cell, const float *perm, const float *decay, int64_t m,
                         int64_t H, int64_t W, void *stream);

/*
 * Committing the rest of a step before the host has read the early-exit flags (no counterpart in
 * the reference, whose host blocks on torch.any() inside the loop, kinetics.py:846).
 *
 * ms_early_exit_verdict evaluates, on the device, what the host does with the flags of a kinetics
 * launch that applied eq_iters_host[q] iterations in part q: *gate = 1 if that launch already equals
 * the reference and met no cell position outside the map, else 0 (the launch has to be repeated
 * with fewer iterations, see `flags` above).
 * ms_step_gate(gate) makes every later call BY THE CALLING THREAD of the entry points that commit
 * results -- ms_scatter_signals[_by_pixel], ms_diffuse_correct, ms_diffuse_molecules,
 * ms_permeate[_by_pixel], ms_degrade_permeate, ms_degrade_molecules -- launch kernels that leave
 * without writing anything if *gate == 0 when they run; ms_step_gate(NULL) ends that.  (The stencil
 * pass writes only scratch and is not gated.)  A host can so enqueue
 *     kinetics, verdict, [gated: write-back, degradation, diffusion, permeation]
 * without a synchronisation in between, read the flags afterwards, and, in the rare case that the
 * verdict was 0, repeat the step from the untouched state.  ms_diffuse_molecules_static ignores it.
 */
int ms_early_exit_verdict(const uint32_t *flags, const int32_t *eq_iters_host, int32_t n_parts,
                          uint32_t *gate, void *stream);
void ms_step_gate(const uint32_t *gate);

/*
 * The small exchanges of a multi-GPU step over NVLink peer memory (csrc/peer.cu; no counterpart in the
 * single-device reference).  `bufs[r]` (r < world <= 16) is the address, in THIS process, of rank r's
 * exchange buffer of ms_peer_buffer_bytes(world, m, W) bytes, zero-initialised before the first call
 * (torch.distributed._symmetric_memory or CUDA IPC provide such mappings).  Every rank makes the same
 * calls in the same order, with `seq` = 1, 2, 3, ... counted separately for the all-reduces and for
 * the halos.  A rank's kernel stores its contribution into every peer's buffer with plain
 * peer-to-peer stores, publishes `seq`, waits for the peers' in its own buffer and reduces: one short
 * launch in stream order instead of a collective call.  If a peer does not show up within ~2 s the
 * kernel sets *status (device word) and returns.
 *   ms_peer_allreduce: `data` (n elements, in place) := sum over ranks (op 0, int64) or max (op 1,
 *     uint32); n * element size <= 16 KB.
 *   ms_peer_halo: the first row of every plane of `map_slab` (m,H,W) goes to the bottom halo of the rank
 *     above, the last row to the top halo of the rank below (circular); on return of the kernel this
 *     rank's own halos of call `seq` are at ms_peer_halo_offset(world, m, W, seq & 1, 0 / 1) bytes
 *     into its buffer: (m, W) floats each, the halo_top / halo_bottom of ms_diffuse_stencil.
 */
size_t ms_peer_buffer_bytes(int32_t world, int64_t m, int64_t W);
size_t ms_peer_halo_offset(int32_t world, int64_t m, int64_t W, int32_t parity, int32_t bottom);
int ms_peer_allreduce(const uint64_t *bufs, int32_t world, int32_t rank, uint64_t seq, void *data, int32_t n,
                      int32_t op, uint32_t *status, void *stream);
int ms_peer_halo(const uint64_t *bufs, int32_t world, int32_t rank, uint64_t seq, const float *map_slab, int64_t m,
                 int64_t H, int64_t W, uint32_t *status, void *stream);

/*
 * World.degrade_molecules()                                    world.py:667-678
 * molecule_map[k] *= decay[k], cell_molecules[:, k] *= decay[k], in place, one launch
 * (decay[k] = exp(-ln 2 / half_life_k), world.py:184).
 */
int ms_degrade_molecules(float *molecule_map, float *cell_molecules, const float *decay, int64_t c,
                         int64_t m, int64_t H, int64_t W, void *stream);

/*
 * degrade_molecules() folded into the next diffuse_molecules(): the stencil
 * multiplies every value it reads (slab and halo rows) by decay[k], the permeation multiplies the
 * cell molecules it reads by decay[k]; with ms_diffuse_correct in between the result is
 * bit-identical to ms_degrade_molecules followed by ms_diffuse_stencil / _correct / ms_permeate,
 * without the extra read and write of the map.
 */
int ms_degrade_diffuse_stencil(const float *map_slab, const float *halo_top, const float *halo_bottom,
                               const float *a, const float *b, const float *decay, int64_t m, int64_t H,
                               int64_t W, float *out, int64_t *sums, void *stream);
int ms_degrade_permeate(const float *cell_molecules_in, float *cell_molecules_out, float *molecule_map,
                        const int32_t *cell_positions, const float *perm, const float *decay, int64_t c,
                        int64_t m, int64_t H, int64_t W, void *stream);

/*
 * ms_diffuse_molecules without scratch from the caller -- no counterpart in the reference.
 * For hosts that cannot allocate device memory around the call (all memory already handed to
 * their own pool, or a step loop that must avoid the device-wide synchronisation that
 * cudaMalloc / cudaFree imply): the stencil output and the plane sums live in a fixed
 * `static __device__` array inside the library, so the call allocates nothing and needs no
 * scratch argument.  Same arithmetic and results as ms_diffuse_molecules.
 *   - Limit: m * H * W <= ms_diffuse_static_max_floats() (4 Mi floats = 16 MiB, e.g. 16 planes
 *     of 512 x 512) and m <= 256.  Larger maps are rejected with MS_ERR_TOO_LARGE and
 *     ms_diffuse_static_last_error() says which limit was hit; nothing is launched.
 *   - Every GPU that loads the library carries the 16 MiB array until the process exits.
 *   - All calls on one GPU share that array; calls that overlap on the device (e.g. from two
 *     streams) race on it.  Its contents survive between calls and are never assumed zero:
 *     each call clears the sums it accumulates into and rewrites the output before reading it.
 *   - The first call on each device looks up the array's device address
 *     (cudaGetSymbolAddress, with that device current); make that first call before any
 *     stream capture begins.
 */
int ms_diffuse_molecules_static(float *molecule_map, const float *a, const float *b, int64_t m,
                                int64_t H, int64_t W, void *stream);
int64_t ms_diffuse_static_max_floats(void);
/* Text of the last MS_ERR_TOO_LARGE refusal of ms_diffuse_molecules_static on the CALLING THREAD
 * (the buffer is thread-local), "" after a call that was not refused. */
const char *ms_diffuse_static_last_error(void);

/*
 * World.diffuse_molecules(), permeation half                   world.py:651-665
 * perm[k] from World._get_permeate (world.py:935-946).  cell_molecules_out may
 * alias cell_molecules_in; molecule_map is updated at the cells' pixels.
 */
int ms_permeate(const float *cell_molecules_in, float *cell_molecules_out, float *molecule_map,
                const int32_t *cell_positions, const float *perm, int64_t c, int64_t m, int64_t H,
                int64_t W, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MAGICSOUP_B200_H */
