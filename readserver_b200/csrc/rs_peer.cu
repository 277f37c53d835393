/* Per-shard counts -> population counts across the GPUs of one box, without a collective library.
 *
 * The reference's front end sums the counts its shard workers send back (src/service/server.cpp:184-197,
 * 404-410).  Here every rank (one process per GPU) owns the reduced counts of the k-mers whose prefix
 * maps to it; K1 adds each finished query's count straight into the owner's accumulation buffer through
 * NVLink peer memory (mode 2 of k_find_interval), so search and reduce-scatter are one kernel.  What is
 * left is a barrier: every rank writes an epoch into its slot of every peer's flag array and waits for
 * all slots of its own.
 *
 * One device allocation per rank, exported through a CUDA IPC handle:
 *   [ flags: RS_MAX_RANKS x u64 | error word | pad to 256 B ][ counts 0 | counts 1 | counts 2: max_queries x u64 each ]
 *   [ scratch: max_queries x u64 (packed queries of the running K1) ]
 * Step e accumulates into buffer e mod 3.  A rank clears the buffer of the NEXT step before it searches, which no
 * peer can touch before it has passed this step's barrier; with three buffers the counts rs_comm_finish(e) returned
 * are still intact after rs_comm_begin(e + 1) (which clears buffer e + 2) and die with rs_comm_begin(e + 2).
 */
#include "rs_internal.h"

#define RS_COMM_HDR 256

struct rs_comm {
  int device = 0;
  uint32_t rank = 0, world = 1;
  uint64_t max_queries = 0;
  uint8_t* base = nullptr;                /* this rank's allocation */
  uint8_t* peer_base[RS_MAX_RANKS] = {};  /* every rank's allocation as mapped here (own = base) */
  uint64_t** d_peer_flags = nullptr;      /* device array of the ranks' flag arrays */
  uint64_t epoch = 0;
  uint64_t step_q = 0;                    /* queries of the step in progress */
  bool connected = false, in_step = false;
  /* the batch whose packed form sits in the scratch area (one pre-pass per step, however many shards) */
  const void* packed_src = nullptr;
  int packed_k = 0, packed_strand = -1;
  void* packed_stream = nullptr;
};

static size_t comm_bytes(uint64_t maxq) { return RS_COMM_HDR + 4 * maxq * sizeof(uint64_t); }
static uint64_t* comm_counts(uint8_t* base, uint64_t maxq, uint64_t epoch) {
  return reinterpret_cast<uint64_t*>(base + RS_COMM_HDR) + (epoch % 3) * maxq;
}

__global__ void k_peer_barrier(uint64_t* const* __restrict__ peer_flags, uint32_t world, uint32_t rank, uint64_t epoch,
                               uint32_t timeout_ms, uint32_t* __restrict__ err) {
  const uint32_t r = threadIdx.x;
  if (r >= world) return;
  __threadfence_system(); /* everything this rank's earlier kernels wrote is visible before the flag */
  volatile uint64_t* theirs = peer_flags[r] + rank;
  *theirs = epoch;
  __threadfence_system();
  volatile uint64_t* mine = peer_flags[rank] + r;
  const long long t0 = clock64();
  const long long limit = (long long)timeout_ms * 2000000ll; /* ~2 GHz */
  while (*mine < epoch) {
    if (clock64() - t0 > limit) {
      *err = 1; /* a peer never arrived: report instead of hanging the device */
      break;
    }
    __nanosleep(200);
  }
  __threadfence_system();
}

cudaError_t rs_launch_peer_barrier(uint64_t* const* peer_flags, uint32_t world, uint32_t rank, uint64_t epoch,
                                   uint32_t timeout_ms, uint32_t* err, cudaStream_t st) {
  k_peer_barrier<<<1, 32, 0, st>>>(peer_flags, world, rank, epoch, timeout_ms, err);
  return cudaGetLastError();
}

int rs_comm_create(int device, int rank, int world, uint64_t max_queries, rs_comm** out) {
  if (!out || rank < 0 || world < 1 || rank >= world || world > RS_MAX_RANKS || max_queries == 0)
    return fail(RS_STATUS_ARG, "bad argument");
  *out = nullptr;
  CU(cudaSetDevice(device));
  rs_comm* c = new rs_comm;
  c->device = device;
  c->rank = (uint32_t)rank;
  c->world = (uint32_t)world;
  c->max_queries = max_queries;
  cudaError_t e = cudaMalloc((void**)&c->base, comm_bytes(max_queries));
  if (e == cudaSuccess) e = cudaMemset(c->base, 0, comm_bytes(max_queries));
  if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_peer_flags, RS_MAX_RANKS * sizeof(uint64_t*));
  if (e == cudaSuccess) e = cudaDeviceSynchronize(); /* the buffers are clear before any handle leaves this process */
  if (e != cudaSuccess) {
    if (c->base) cudaFree(c->base);
    delete c;
    return fail(RS_STATUS_CUDA, "rs_comm_create: %s", cudaGetErrorString(e));
  }
  c->peer_base[rank] = c->base;
  *out = c;
  return RS_STATUS_OK;
}

int rs_comm_export(rs_comm* c, uint8_t handle[RS_COMM_HANDLE_BYTES]) {
  if (!c || !handle) return fail(RS_STATUS_ARG, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == RS_COMM_HANDLE_BYTES, "IPC handle size");
  CU(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, c->base));
  memcpy(handle, &h, sizeof(h));
  return RS_STATUS_OK;
}

int rs_comm_connect(rs_comm* c, const uint8_t* handles) {
  if (!c || (!handles && c->world > 1)) return fail(RS_STATUS_ARG, "null argument");
  if (c->connected) return fail(RS_STATUS_ARG, "already connected");
  CU(cudaSetDevice(c->device));
  for (uint32_t r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + (size_t)r * RS_COMM_HANDLE_BYTES, sizeof(h));
    void* p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer_base[r] = (uint8_t*)p;
  }
  uint64_t* flags[RS_MAX_RANKS] = {};
  for (uint32_t r = 0; r < c->world; ++r) flags[r] = reinterpret_cast<uint64_t*>(c->peer_base[r]);
  CU(cudaMemcpy(c->d_peer_flags, flags, sizeof(flags), cudaMemcpyHostToDevice));
  c->connected = true;
  return RS_STATUS_OK;
}

void rs_comm_destroy(rs_comm* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (uint32_t r = 0; r < c->world; ++r)
    if (r != c->rank && c->peer_base[r]) cudaIpcCloseMemHandle(c->peer_base[r]);
  if (c->d_peer_flags) cudaFree(c->d_peer_flags);
  if (c->base) cudaFree(c->base);
  delete c;
}

int rs_comm_begin(rs_comm* c, uint64_t Q, void* stream) {
  if (!c || !c->connected) return fail(RS_STATUS_ARG, "communicator not connected");
  if (c->in_step) return fail(RS_STATUS_ARG, "previous step not finished");
  if (Q > c->max_queries) return fail(RS_STATUS_ARG, "batch larger than the communicator's capacity");
  CU(cudaSetDevice(c->device));
  ++c->epoch;
  c->step_q = Q;
  c->in_step = true;
  c->packed_src = nullptr;
  /* clear the NEXT step's buffer: peers add into it only after they passed this step's barrier, which
   * needs this rank's flag, which is written after this memset in stream order */
  CU(cudaMemsetAsync(comm_counts(c->base, c->max_queries, c->epoch + 1), 0, c->max_queries * sizeof(uint64_t),
                     (cudaStream_t)stream));
  return RS_STATUS_OK;
}

int rs_count_accumulate_dev(const rs_index* ix, rs_comm* c, const void* d_kmers, int k, size_t Q, int strand,
                            void* stream) {
  if (!ix || !c || (Q && !d_kmers)) return fail(RS_STATUS_ARG, "null argument");
  if (!c->in_step || Q != c->step_q) return fail(RS_STATUS_ARG, "rs_comm_begin was not called for this batch size");
  if (k < 1 || k > 31) return fail(RS_STATUS_ARG, "the fused count supports fixed-length queries of 1..31 characters");
  if (strand != 0 && strand != 1) return fail(RS_STATUS_ARG, "strand must be 0 or 1");
  if (ix->device != c->device) return fail(RS_STATUS_ARG, "index and communicator live on different devices");
  rs_peers peers;
  memset(&peers, 0, sizeof(peers));
  peers.world = c->world;
  peers.rank = c->rank;
  for (uint32_t r = 0; r < c->world; ++r) peers.buf[r] = comm_counts(c->peer_base[r], c->max_queries, c->epoch);
  uint64_t* scratch = reinterpret_cast<uint64_t*>(c->base + RS_COMM_HDR) + 3 * c->max_queries;
  const bool reuse = c->packed_src == d_kmers && c->packed_k == k && c->packed_strand == strand && c->packed_stream == stream;
  CU(rs_launch_find_interval(ix->dev, (const uint8_t*)d_kmers, nullptr, (uint32_t)k, Q, scratch, nullptr, reuse ? 3 : 2, strand,
                             (unsigned long long*)ix->d_stats, (cudaStream_t)stream, &peers));
  c->packed_src = d_kmers;
  c->packed_k = k;
  c->packed_strand = strand;
  c->packed_stream = stream;
  return RS_STATUS_OK;
}

int rs_comm_finish(rs_comm* c, void* stream, void** d_counts) {
  if (!c || !c->in_step) return fail(RS_STATUS_ARG, "no step in progress");
  CU(cudaSetDevice(c->device));
  uint32_t* err = reinterpret_cast<uint32_t*>(c->base + RS_MAX_RANKS * sizeof(uint64_t));
  CU(rs_launch_peer_barrier(c->d_peer_flags, c->world, c->rank, c->epoch, 5000, err, (cudaStream_t)stream));
  c->in_step = false;
  if (d_counts) *d_counts = comm_counts(c->base, c->max_queries, c->epoch);
  return RS_STATUS_OK;
}

int rs_comm_error(rs_comm* c, int* timed_out) {
  if (!c || !timed_out) return fail(RS_STATUS_ARG, "null argument");
  CU(cudaSetDevice(c->device));
  uint32_t v = 0;
  CU(cudaMemcpy(&v, c->base + RS_MAX_RANKS * sizeof(uint64_t), sizeof(v), cudaMemcpyDeviceToHost));
  *timed_out = (int)v;
  return RS_STATUS_OK;
}
