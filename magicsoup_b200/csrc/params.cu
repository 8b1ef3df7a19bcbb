// Kinetics.set_cell_params on the device (reference python/magicsoup/kinetics.py:521-625): domain
// tokens -> the nine parameter tensors, written straight into the parameter rows.
//
// The reference gathers (c,p,d) and (c,p,d,s) temporaries from its token maps and reduces them over
// the domain axis with einsum / nanmean / sum; here one warp takes one (cell, protein) row, walks
// the protein's few domains and lets its lanes own the signals, so nothing but the row is written.
// Integer results (N, Nf, Nb, A) are exact; float results follow the reference's formulas in
// float32 (means as sum / count in domain order, powf, expf), the energy sum in float64.
#include <math.h>

#include "ms_common.cuh"

namespace ms {

constexpr int ROW_WARPS = 8;

__global__ void __launch_bounds__(32 * ROW_WARPS) set_cell_params_kernel(
    const ms_token_tables T, const float *__restrict__ energies, float abs_temp, const int *__restrict__ rows,
    const int *__restrict__ prot_offsets, const int *__restrict__ dom_offsets, const int *__restrict__ doms,
    long long n_cells, const ms_kinetics_params_out out) {
  const int lane = threadIdx.x & 31;
  const long long w = (long long)blockIdx.x * ROW_WARPS + (threadIdx.x >> 5);
  const int p = (int)out.p, s = (int)out.s;
  if (w >= n_cells * p) return;
  const long long k = w / p;
  const int j = (int)(w - k * p);
  const long long row = (long long)rows[k] * p + j;  // (cell row, protein) in the (c,p[,s]) tensors
  const int po = prot_offsets[k], n_prot = prot_offsets[k + 1] - po;
  int d0 = 0, d1 = 0;
  if (j < n_prot) {
    d0 = dom_offsets[po + j];
    d1 = dom_offsets[po + j + 1];
  }
  const float qnan = __int_as_float(0x7fc00000);

  // per-protein scalars (every lane computes them; a protein has a handful of domains)
  float vsum = 0.0f, ksum = 0.0f;
  int vcnt = 0, kcnt = 0;
  for (int d = d0; d < d1; d++) {
    const int dt = doms[5 * d], i0 = doms[5 * d + 1], i1 = doms[5 * d + 2];
    const bool reg = dt == 3;
    const float vm = T.vmax_weights[reg ? 0 : i0];  // kinetics.py:560 (undefined entries are NaN)
    const float km = reg ? qnan : T.km_weights[i1]; // kinetics.py:562, 608
    if (!isnan(vm)) { vsum = vsum + vm; vcnt++; }
    if (!isnan(km)) { ksum = ksum + km; kcnt++; }
  }
  const float vmax = vcnt ? vsum / (float)vcnt : 0.0f;  // nanmean(...).nan_to_num(0), kinetics.py:573
  const float kmn = kcnt ? ksum / (float)kcnt : 0.0f;   // kinetics.py:608

  double e_part = 0.0;
  for (int i = lane; i < s; i += 32) {
    int n = 0, nf = 0, nb = 0;
    float af = 0.0f, rsum = 0.0f;
    int rcnt = 0;
    for (int d = d0; d < d1; d++) {
      const int dt = doms[5 * d], i0 = doms[5 * d + 1], i1 = doms[5 * d + 2], i2 = doms[5 * d + 3], i3 = doms[5 * d + 4];
      const int sign = T.signs[i2];
      // vectors of the domain's own type; token 0 (all other types) maps to the zero vector
      const int react = T.reaction_M[(long long)(dt == 1 ? i3 : 0) * s + i];
      const int trnsp = T.transport_M[(long long)(dt == 2 ? i3 : 0) * s + i];
      const int eff = T.effector_M[(long long)(dt == 3 ? i3 : 0) * s + i];
      const int nd = (react + trnsp) * sign;  // kinetics.py:597
      n += nd;
      nf += nd < 0 ? -nd : 0;                 // kinetics.py:603-604
      nb += nd > 0 ? nd : 0;
      const int hill = T.hill_numbers[dt == 3 ? i0 : 0];
      af = af + (float)eff * (float)(sign * hill);  // kinetics.py:578-580
      const float kv = (float)eff * (dt == 3 ? T.km_weights[i1] : qnan);  // kinetics.py:584-585
      if (kv != 0.0f && !isnan(kv)) { rsum = rsum + kv; rcnt++; }          // :588-589
    }
    const int a = (int)af;
    const float kmr = rcnt ? rsum / (float)rcnt : 0.0f;
    out.N[row * s + i] = n;
    out.Nf[row * s + i] = nf;
    out.Nb[row * s + i] = nb;
    out.A[row * s + i] = a;
    out.Kmr[row * s + i] = powf(kmr, (float)a);  // kinetics.py:592 (0^0 = 1 where nothing regulates)
    e_part += (double)n * (double)energies[i];
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) e_part += __shfl_xor_sync(0xffffffffu, e_part, d);
  if (lane == 0) {
    const float E = (float)e_part;                                   // kinetics.py:612
    float ke = expf((-E / abs_temp) / (float)8.31446261815324);       // :613
    ke = fminf(fmaxf(ke, V_EPS), V_MAX);                              // clamp passes NaN through like torch
    if (isnan(E)) ke = qnan;
    const bool fwd = ke >= 1.0f;
    float kmf = fwd ? kmn : kmn / ke, kmb = fwd ? kmn * ke : kmn;     // :623-625
    if (!isnan(kmf)) kmf = fminf(fmaxf(kmf, V_EPS), V_MAX);
    if (!isnan(kmb)) kmb = fminf(fmaxf(kmb, V_EPS), V_MAX);
    out.Vmax[row] = vmax;
    out.Ke[row] = ke;
    out.Kmf[row] = kmf;
    out.Kmb[row] = kmb;
  }
}

}  // namespace ms

extern "C" int ms_set_cell_params(const ms_token_tables *tables, const float *mol_energies, float abs_temp,
                                  const int32_t *rows, const int32_t *prot_offsets, const int32_t *dom_offsets,
                                  const int32_t *doms, int64_t n_cells, const ms_kinetics_params_out *out,
                                  void *stream) {
  if (!tables || !out || n_cells < 0 || out->p < 0 || out->s < 0) return MS_ERR_INVALID_ARGUMENT;
  if (n_cells == 0 || out->p == 0) return MS_OK;
  if (!rows || !prot_offsets || !dom_offsets || !mol_energies || !tables->vmax_weights || !tables->km_weights ||
      !tables->signs || !tables->hill_numbers || !tables->reaction_M || !tables->transport_M || !tables->effector_M ||
      !out->N || !out->Nf || !out->Nb || !out->A || !out->Kmr || !out->Ke || !out->Kmf || !out->Kmb || !out->Vmax)
    return MS_ERR_INVALID_ARGUMENT;
  const long long n_rows = n_cells * out->p;
  const unsigned blocks = (unsigned)((n_rows + ms::ROW_WARPS - 1) / ms::ROW_WARPS);
  ms::set_cell_params_kernel<<<blocks, 32 * ms::ROW_WARPS, 0, (cudaStream_t)stream>>>(
      *tables, mol_energies, abs_temp, rows, prot_offsets, dom_offsets, doms, n_cells, *out);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}
