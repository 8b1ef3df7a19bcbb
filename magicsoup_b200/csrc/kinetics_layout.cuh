// Region layout of the lane path (kinetics_lanes.cuh), shared by host and device code.
#pragma once

namespace ms {

__host__ __device__ inline int round32(int x) { return (x + 31) & ~31; }
__host__ __device__ inline int pow2_at_least_32(int x) {
  int r = 32;
  while (r < x) r <<= 1;
  return r;
}

struct LaneLayout {
  int S2, PA, PA2, TS, EX, RG, CH, R;
  // static arrays (cache record), word offsets in the warp's region
  int oCol, oKe, oKmf, oKmb, oVmax, oPf, oEnt8, oExt8, oRare, oRegw, oRegk, oCsc, static_words;
  // working arrays
  int oX0, oX1, oRegf, oNv, oT, oV, oF, oFk, words;
  int oQueue, q_cap;  // scan-time element queue (2 words each): the last 2*q_cap words
};

constexpr int LANE_MAX_PS = 64;   // p and s limit of the lane path
constexpr int LANE_EXT_MAX = 7;   // extension chunks per protein (<= 32 entries per side)
constexpr int LANE_REG_MAX = 15;  // regulatory entries per protein

__host__ __device__ inline LaneLayout make_lane_layout(int p, int s) {
  LaneLayout L;
  L.S2 = pow2_at_least_32(s);
  L.PA = round32(p);
  L.PA2 = L.PA + 32;  // + the dummy protein of the padding entries
  L.R = 31;
  L.TS = 3 * L.S2 + 32;
  L.EX = 16;
  L.RG = L.PA;
  L.CH = ((7 * p / 2 + 3) / 4 + s / 2 + 3) & ~3;
  L.q_cap = round32((7 * p + 1) / 2 < 64 ? 64 : (7 * p + 1) / 2);
  int o = 0;
  L.oCol = o; o += L.S2;
  L.oKe = o; o += L.PA;
  L.oKmf = o; o += L.PA;
  L.oKmb = o; o += L.PA;
  L.oVmax = o; o += L.PA;
  L.oPf = o; o += L.PA;
  L.oEnt8 = o; o += 2 * L.PA;
  L.oExt8 = o; o += 2 * L.EX;
  L.oRare = o; o += 32;
  L.oRegw = o; o += L.RG;
  L.oRegk = o; o += L.RG;
  L.oCsc = o; o += 2 * L.CH;
  L.static_words = o;
  L.oX0 = o; o += L.S2;
  L.oX1 = o; o += L.S2;
  L.oRegf = o; o += L.RG;
  L.oNv = o; o += 4 * L.CH;
  L.oT = o; o += L.TS;
  L.oV = o; o += L.PA2;
  L.oF = o; o += L.PA2;
  L.oFk = o; o += L.PA2;
  // the queue and the scan's staging slots (8 x 512 B) share the region while the tile is scanned
  if (o < 1024 + 2 * L.q_cap) o = 1024 + 2 * L.q_cap;
  L.words = (o + 31) & ~31;
  {  // the queue gets whatever the region leaves behind the staging slots and behind the static
     // arrays + the list builder's temporaries (3*PA + 2*S2 words from oX0): costs nothing
    const int busy = L.static_words + 3 * L.PA + 2 * L.S2;
    L.q_cap = (L.words - (busy > 1024 ? busy : 1024)) / 2;
  }
  L.oQueue = L.words - 2 * L.q_cap;
  return L;
}

}  // namespace ms
