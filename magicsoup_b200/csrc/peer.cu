// The small exchanges of a multi-GPU step over NVLink peer memory (no counterpart in the single-device
// reference; magicsoup_b200/distributed.py): the all-reduce of the 32 early-exit flag words, the
// all-reduce of the 5*m exact plane sums of the mass correction, and the two halo rows of the
// diffusion stencil.  Each is a few hundred bytes to 128 KB; as NCCL calls they cost 20-60 us apiece
// on the device, in a chain every step waits for.  Here a rank's kernel STORES its contribution into
// every peer's buffer (plain peer-to-peer stores through NVSwitch), publishes a sequence number with
// system-scope release, waits for the peers' sequence numbers in its OWN memory and reduces what has
// arrived -- one short kernel per exchange, launched in stream order like any other kernel of the step.
//
// Every rank owns one buffer of ms_peer_buffer_bytes(); `bufs[r]` is rank r's buffer as mapped into
// this process (torch.distributed._symmetric_memory hands those out).  Layout:
//   [0, 256)        all-reduce sequence numbers   [2 parities][16 ranks] uint64
//   [256, 512)      halo sequence numbers         [2 parities][2: top, bottom] uint64
//   [512, ...)      all-reduce slots              [2 parities][world][SLOT_BYTES]
//   then            halo rows                     [2 parities][2: top, bottom][m * W floats, 256-byte aligned]
// Two parities: call k + 2 reuses the slots of call k, and a rank can only get there after every
// peer has published call k + 1, i.e. after every peer has finished reading call k.
// A wait that sees nothing for ~2 s sets *status and gives up (the results are then garbage; the
// host raises): a missing peer never hangs the GPU.
#include "ms_common.cuh"

namespace ms {

constexpr int PEER_MAX_RANKS = 16;
constexpr size_t PEER_SLOT_BYTES = 16 * 1024;
constexpr size_t PEER_SLOTS_OFFSET = 512;
constexpr unsigned long long PEER_SPIN_LIMIT = 20ull * 1000 * 1000;  // x 100 ns

struct PeerBufs {
  unsigned long long base[PEER_MAX_RANKS];
};

__host__ __device__ inline size_t peer_halo_row_bytes(long long m, long long W) {
  return ((size_t)m * (size_t)W * sizeof(float) + 255) & ~(size_t)255;
}
__host__ __device__ inline size_t peer_halo_offset(int world) {
  return PEER_SLOTS_OFFSET + 2 * (size_t)world * PEER_SLOT_BYTES;
}

__device__ __forceinline__ void publish(unsigned long long *flag, unsigned long long seq) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(seq) : "memory");
}
__device__ __forceinline__ unsigned long long observe(const unsigned long long *flag) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
  return v;
}
// false on timeout
__device__ __forceinline__ bool wait_for(const unsigned long long *flag, unsigned long long seq) {
  for (unsigned long long spins = 0; observe(flag) < seq; spins++) {
    if (spins > PEER_SPIN_LIMIT) return false;
    __nanosleep(100);
  }
  return true;
}

// OP 0: int64 sum, OP 1: uint32 max.  One CTA.
template <int OP>
__global__ void __launch_bounds__(256) peer_allreduce_kernel(const PeerBufs bufs, int world, int rank,
                                                             unsigned long long seq, void *data, int n,
                                                             unsigned *status) {
  const int parity = (int)(seq & 1ull);
  const size_t mine = PEER_SLOTS_OFFSET + ((size_t)parity * world + rank) * PEER_SLOT_BYTES;
  for (int r = 0; r < world; r++) {  // my contribution into every rank's slot for me (my own included)
    char *dst = reinterpret_cast<char *>(bufs.base[r]) + mine;
    if (OP == 0) {
      for (int i = threadIdx.x; i < n; i += blockDim.x)
        reinterpret_cast<long long *>(dst)[i] = reinterpret_cast<const long long *>(data)[i];
    } else {
      for (int i = threadIdx.x; i < n; i += blockDim.x)
        reinterpret_cast<unsigned *>(dst)[i] = reinterpret_cast<const unsigned *>(data)[i];
    }
  }
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < world)
    publish(reinterpret_cast<unsigned long long *>(bufs.base[threadIdx.x]) + parity * PEER_MAX_RANKS + rank, seq);
  if ((int)threadIdx.x < world) {
    const unsigned long long *flag =
        reinterpret_cast<const unsigned long long *>(bufs.base[rank]) + parity * PEER_MAX_RANKS + threadIdx.x;
    if (!wait_for(flag, seq)) atomicOr(status, 1u);
  }
  __syncthreads();
  const char *slots = reinterpret_cast<const char *>(bufs.base[rank]) + PEER_SLOTS_OFFSET +
                      (size_t)parity * world * PEER_SLOT_BYTES;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {  // what the peers stored: read past the L1 of this SM
    if (OP == 0) {
      long long acc = 0;
      for (int r = 0; r < world; r++)
        acc += __ldcv(reinterpret_cast<const long long *>(slots + (size_t)r * PEER_SLOT_BYTES) + i);
      reinterpret_cast<long long *>(data)[i] = acc;
    } else {
      unsigned acc = 0;
      for (int r = 0; r < world; r++) {
        const unsigned v = __ldcv(reinterpret_cast<const unsigned *>(slots + (size_t)r * PEER_SLOT_BYTES) + i);
        acc = v > acc ? v : acc;
      }
      reinterpret_cast<unsigned *>(data)[i] = acc;
    }
  }
}

// my first row of every plane -> the BOTTOM halo of the rank above, my last row -> the TOP halo of the
// rank below (circular); then wait until my own two halos of this call have arrived.  One CTA.
__global__ void __launch_bounds__(1024) peer_halo_kernel(const PeerBufs bufs, int world, int rank, unsigned long long seq,
                                                         const float *__restrict__ map, int m, int H, int W,
                                                         unsigned *status) {
  const int parity = (int)(seq & 1ull);
  const int above = (rank + world - 1) % world, below = (rank + 1) % world;
  const size_t row_bytes = peer_halo_row_bytes(m, W), base = peer_halo_offset(world) + (size_t)parity * 2 * row_bytes;
  float *to_above = reinterpret_cast<float *>(reinterpret_cast<char *>(bufs.base[above]) + base + row_bytes);  // its bottom
  float *to_below = reinterpret_cast<float *>(reinterpret_cast<char *>(bufs.base[below]) + base);              // its top
  const size_t plane = (size_t)H * W;
  for (int i = threadIdx.x; i < m * W; i += blockDim.x) {
    const int k = i / W, y = i - k * W;
    to_above[i] = map[k * plane + y];
    to_below[i] = map[k * plane + (size_t)(H - 1) * W + y];
  }
  __threadfence_system();
  __syncthreads();
  unsigned long long *flags_above = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(bufs.base[above]) + 256);
  unsigned long long *flags_below = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(bufs.base[below]) + 256);
  if (threadIdx.x == 0) publish(flags_above + parity * 2 + 1, seq);  // its bottom halo is there
  if (threadIdx.x == 1) publish(flags_below + parity * 2 + 0, seq);  // its top halo is there
  if (threadIdx.x < 2) {
    const unsigned long long *flag =
        reinterpret_cast<const unsigned long long *>(reinterpret_cast<const char *>(bufs.base[rank]) + 256) + parity * 2 + threadIdx.x;
    if (!wait_for(flag, seq)) atomicOr(status, 1u);
  }
}

}  // namespace ms

using namespace ms;

static bool fill_bufs(PeerBufs &b, const uint64_t *bufs, int32_t world, int32_t rank) {
  if (!bufs || world < 1 || world > PEER_MAX_RANKS || rank < 0 || rank >= world) return false;
  for (int r = 0; r < world; r++) {
    if (!bufs[r]) return false;
    b.base[r] = bufs[r];
  }
  return true;
}

extern "C" size_t ms_peer_buffer_bytes(int32_t world, int64_t m, int64_t W) {
  if (world < 1 || world > PEER_MAX_RANKS || m < 0 || W < 0) return 0;
  return peer_halo_offset(world) + 4 * peer_halo_row_bytes(m, W);
}

extern "C" size_t ms_peer_halo_offset(int32_t world, int64_t m, int64_t W, int32_t parity, int32_t bottom) {
  return peer_halo_offset(world) + ((size_t)(parity & 1) * 2 + (bottom ? 1 : 0)) * peer_halo_row_bytes(m, W);
}

extern "C" int ms_peer_allreduce(const uint64_t *bufs, int32_t world, int32_t rank, uint64_t seq, void *data, int32_t n,
                                 int32_t op, uint32_t *status, void *stream) {
  PeerBufs b;
  if (!fill_bufs(b, bufs, world, rank) || !data || !status || n < 0 || seq == 0 || (op != 0 && op != 1))
    return MS_ERR_INVALID_ARGUMENT;
  if ((size_t)n * (op == 0 ? 8 : 4) > PEER_SLOT_BYTES) return MS_ERR_TOO_LARGE;
  if (n == 0) return MS_OK;
  if (op == 0)
    peer_allreduce_kernel<0><<<1, 256, 0, (cudaStream_t)stream>>>(b, world, rank, seq, data, n, status);
  else
    peer_allreduce_kernel<1><<<1, 256, 0, (cudaStream_t)stream>>>(b, world, rank, seq, data, n, status);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}

extern "C" int ms_peer_halo(const uint64_t *bufs, int32_t world, int32_t rank, uint64_t seq, const float *map_slab,
                            int64_t m, int64_t H, int64_t W, uint32_t *status, void *stream) {
  PeerBufs b;
  if (!fill_bufs(b, bufs, world, rank) || !map_slab || !status || seq == 0 || m < 1 || H < 1 || W < 1 ||
      m * W > INT32_MAX || H > INT32_MAX)
    return MS_ERR_INVALID_ARGUMENT;
  peer_halo_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(b, world, rank, seq, map_slab, (int)m, (int)H, (int)W, status);
  MS_CUDA_CHECK(cudaGetLastError());
  return MS_OK;
}
