// prefix_index.cu — the KV router's prefix index with batched GPU lookups (SURVEY §8f "next-1").
//
// Reference: RadixTree in lib/llm/src/kv_router/indexer.rs:187-387 — a tree of blocks keyed by LocalBlockHash whose
// nodes carry the set of workers holding that block; `find_matches` (:239-277) walks a request's block hashes from the
// root and counts, per worker, how many leading blocks it holds; `apply_event` (:284-378) inserts / removes blocks per
// worker through a per-worker table keyed by the engine's ExternalSequenceBlockHash; `remove_worker` (:380-386).
// There it is a single-threaded actor over Rc<RefCell<..>> nodes, one request at a time.
//
// Here the structure is split by what each side is good at:
//   * event application stays on the host, serial and in event order (it IS order-dependent), in a plain C++ mirror of
//     the reference's semantics — including blocks shared between parents through the per-worker table and the
//     "last worker left: drop all children" rule;
//   * every change is mirrored into ONE flat device array: an open-addressing edge table keyed by (parent node, parent
//     generation, tokens hash) whose 32-byte records also carry the child's id, worker bitset and generation.
//     Dropping a node's children is a generation bump (stale edges simply stop matching);
//   * find_matches runs for a whole batch of requests in one launch, one thread per request, each step ONE 32-byte
//     dependent random read; per-worker scores are kept as bit-sliced counters (one 64-bit plane per
//     binary digit), so adding a node's worker set costs a ripple of AND/XORs instead of a 64-way loop.
// Worker sets: up to 64 workers fit the 64-bit word an edge record carries (one dependent read per step). An index
// created for more (up to 512) keeps ceil(W/64) words per node in a second, node-indexed device table: the edge then
// carries the child's worker COUNT (for the early exit), one thread walks a request once per 64-worker word (grid.y)
// and reads that word of the child's set next to the edge it follows — two reads per step that do not depend on each
// other. The reference's set is an unbounded HashSet<WorkerId> (indexer.rs:163-170); the host layer maps WorkerId -> slot.
// Nodes are reference-counted by their incoming parent edges and per-worker lookup entries, as the reference's
// Rc<RefCell<RadixBlock>> are: a subtree dropped by "last worker left" is reclaimed and its ids are reused.
#include <string.h>

#include <algorithm>
#include <deque>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "common.cuh"

namespace kvdp {
namespace {

// 32 bytes = one sector per probe. The child's worker set and generation ride along in the edge, so a walk step is
// ONE dependent random read instead of two (edge, then node record); the host rewrites the copies whenever a node's
// set or generation changes (it keeps, per node, the table slots of its incoming edges).
struct EdgeRec {
  uint64_t hash;
  uint32_t parent;
  uint32_t gen;            // parent's generation when the edge was inserted
  uint64_t child_workers;
  uint32_t child;
  uint32_t child_gen1;     // child's generation + 1; 0 marks an empty slot
};

__host__ __device__ __forceinline__ uint64_t edge_slot_hash(uint64_t hash, uint32_t parent, uint32_t gen) {
  uint64_t x = hash ^ (((uint64_t)parent << 32) | gen) * 0x9E3779B97F4A7C15ULL;
  x ^= x >> 32;
  x *= 0xD6E8FEB86659FD93ULL;
  x ^= x >> 32;
  return x;
}

constexpr int kPlanes = 16;  // bit-sliced counters: depths up to 65535 blocks
constexpr int kMaxWorkerWords = 8;  // 512 workers per index

// MULTI = false: worker sets of <= 64 workers ride in the edge record. MULTI = true: blockIdx.y selects a 64-worker word
// of the node-indexed set table `wsets` ([node][words]); the edge's worker field then holds the child's worker count.
template <bool MULTI>
__global__ void __launch_bounds__(128) kvdp_index_find_kernel(const EdgeRec* __restrict__ edges, uint64_t mask,
                                                              uint32_t root_gen,
                                                              const uint64_t* __restrict__ hashes,
                                                              const int64_t* __restrict__ seq_offsets, int64_t n_req,
                                                              int early_exit, int num_workers,
                                                              uint32_t* __restrict__ scores,
                                                              int32_t* __restrict__ matched_depth,
                                                              const uint64_t* __restrict__ wsets, int words) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_req) return;
  const int word = MULTI ? (int)blockIdx.y : 0;
  const int64_t s0 = seq_offsets[r], s1 = seq_offsets[r + 1];
  uint64_t plane[kPlanes];
#pragma unroll
  for (int p = 0; p < kPlanes; ++p) plane[p] = 0;
  uint32_t node = 0, gen = root_gen;
  int depth = 0;
  for (int64_t i = s0; i < s1; ++i) {
    const uint64_t h = __ldg(hashes + i);
    uint64_t slot = edge_slot_hash(h, node, gen) & mask;
    bool found = false;
    uint4 a, b;
    for (;;) {
      a = __ldg(reinterpret_cast<const uint4*>(edges + slot));       // hash, parent, gen
      b = __ldg(reinterpret_cast<const uint4*>(edges + slot) + 1);   // child_workers, child, child_gen + 1
      if (b.w == 0) break;                                            // empty slot: not present
      const uint64_t eh = (uint64_t)a.x | ((uint64_t)a.y << 32);
      if (eh == h && a.z == node && a.w == gen) {
        found = true;
        break;
      }
      slot = (slot + 1) & mask;
    }
    if (!found) break;
    uint64_t carry;
    int nworkers;
    if (MULTI) {
      carry = __ldg(wsets + (size_t)b.z * (size_t)words + word);     // independent of the next edge read
      nworkers = (int)b.x;
    } else {
      carry = (uint64_t)b.x | ((uint64_t)b.y << 32);
      nworkers = __popcll(carry);
    }
    ++depth;
#pragma unroll
    for (int p = 0; p < kPlanes; ++p) {   // planes += worker mask
      const uint64_t t = plane[p] & carry;
      plane[p] ^= carry;
      carry = t;
    }
    if (early_exit && nworkers == 1) break;
    node = b.z;
    gen = b.w - 1;
  }
  if (matched_depth && word == 0) matched_depth[r] = depth;
  const int w_end = min(num_workers - word * 64, 64);
  for (int w = 0; w < w_end; ++w) {
    uint32_t c = 0;
#pragma unroll
    for (int p = 0; p < kPlanes; ++p) c |= (uint32_t)((plane[p] >> w) & 1ULL) << p;
    scores[r * (int64_t)num_workers + word * 64 + w] = c;
  }
}

// table[idx[i]][0..words) = recs[i][0..words): worker-set words of the nodes whose sets changed
__global__ void kvdp_index_scatter_words_kernel(uint64_t* __restrict__ table, const uint32_t* __restrict__ idx,
                                                const uint64_t* __restrict__ recs, int n, int words) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n * words) table[(size_t)idx[i / words] * words + i % words] = recs[i];
}

template <typename Rec>
__global__ void kvdp_index_scatter_kernel(Rec* __restrict__ table, const uint32_t* __restrict__ idx,
                                          const Rec* __restrict__ recs, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) table[idx[i]] = recs[i];
}

struct HostNode {
  std::unordered_map<uint64_t, uint32_t> children;  // tokens hash -> node id
  std::vector<uint32_t> in_slots;                    // edge-table slots of edges that point at this node
  uint64_t workers[kMaxWorkerWords] = {};
  uint32_t gen = 0;
  uint32_t refs = 0;  // parent edges + per-worker lookup entries that name this node (the reference's Rc count)
  std::deque<int64_t> recent_uses;  // access times (ns) inside the expiration window; only kept when tracking is on

  bool has(int w) const { return workers[w >> 6] >> (w & 63) & 1ULL; }
  void add(int w) { workers[w >> 6] |= 1ULL << (w & 63); }
  void drop(int w) { workers[w >> 6] &= ~(1ULL << (w & 63)); }
  int count() const {
    int c = 0;
    for (int k = 0; k < kMaxWorkerWords; ++k) c += __builtin_popcountll(workers[k]);
    return c;
  }
  bool none() const {
    for (int k = 0; k < kMaxWorkerWords; ++k)
      if (workers[k]) return false;
    return true;
  }
};

}  // namespace
}  // namespace kvdp

using namespace kvdp;

struct kvdp_index {
  int device = -1;
  int max_workers = 64;
  int64_t max_nodes = 0;
  uint64_t capacity = 0;  // edge table slots (power of two)
  uint64_t live_slots = 0;
  int words = 1;          // 64-bit words per worker set (1: sets ride in the edge records)
  std::vector<HostNode> nodes;
  std::vector<uint32_t> free_nodes;  // reclaimed node ids
  std::vector<std::unordered_map<uint64_t, uint32_t>> lookup;  // [worker slot]: engine block hash -> node
  std::vector<uint8_t> lookup_present;
  std::vector<int64_t> member_nodes;  // nodes whose worker set holds this slot (0 after remove_worker = slot reusable)
  uint64_t* d_wsets = nullptr;        // words > 1: [max_nodes][words]
  std::vector<uint32_t> dirty_nodes;
  std::vector<uint8_t> node_dirty_flag;
  std::vector<EdgeRec> h_edges;
  EdgeRec* d_edges = nullptr;
  std::vector<uint32_t> dirty_edges;
  std::vector<uint8_t> edge_dirty_flag;
  // upload staging (grown on demand)
  uint32_t* d_idx = nullptr;
  void* d_recs = nullptr;
  size_t stage_cap = 0;
  cudaStream_t stream = nullptr;
  bool full_upload_needed = true;
  std::recursive_mutex mu;  // events may arrive from the publisher thread while another thread scores requests
  // host-form lookups: one pinned + one device staging buffer, grown on demand
  char* h_io = nullptr;
  char* d_io = nullptr;
  size_t io_cap = 0;
};

namespace {

void mark_edge(kvdp_index* ix, uint64_t slot) {
  if (!ix->edge_dirty_flag[slot]) {
    ix->edge_dirty_flag[slot] = 1;
    ix->dirty_edges.push_back((uint32_t)slot);
  }
}

// A node's worker set or generation changed: refresh the copies carried by its incoming edges.
uint64_t edge_worker_field(const kvdp_index* ix, const HostNode& h) {
  return ix->words == 1 ? h.workers[0] : (uint64_t)h.count();
}

void mark_node(kvdp_index* ix, uint32_t n) {
  const HostNode& h = ix->nodes[n];
  for (uint32_t slot : h.in_slots) {
    EdgeRec& e = ix->h_edges[slot];
    if (e.child_gen1 != 0 && e.child == n) {
      e.child_workers = edge_worker_field(ix, h);
      e.child_gen1 = h.gen + 1;
      mark_edge(ix, slot);
    }
  }
  if (ix->words > 1 && !ix->node_dirty_flag[n]) {
    ix->node_dirty_flag[n] = 1;
    ix->dirty_nodes.push_back(n);
  }
}

int insert_edge(kvdp_index* ix, uint32_t parent, uint64_t hash, uint32_t child) {
  const uint32_t gen = ix->nodes[parent].gen;
  const uint64_t mask = ix->capacity - 1;
  uint64_t slot = edge_slot_hash(hash, parent, gen) & mask;
  for (uint64_t probes = 0; probes < ix->capacity; ++probes) {
    EdgeRec& e = ix->h_edges[slot];
    if (e.child_gen1 == 0) {
      e.hash = hash;
      e.parent = parent;
      e.gen = gen;
      e.child = child;
      e.child_workers = edge_worker_field(ix, ix->nodes[child]);
      e.child_gen1 = ix->nodes[child].gen + 1;
      ++ix->live_slots;
      ix->nodes[child].in_slots.push_back((uint32_t)slot);
      mark_edge(ix, slot);
      return 0;
    }
    slot = (slot + 1) & mask;
  }
  return fail(KVDP_ERANGE, "prefix index: edge table full");
}

// Rebuild the edge table from the authoritative host maps (drops the dead generations).
int rebuild_edges(kvdp_index* ix) {
  std::fill(ix->h_edges.begin(), ix->h_edges.end(), EdgeRec{});
  std::fill(ix->edge_dirty_flag.begin(), ix->edge_dirty_flag.end(), 0);
  ix->dirty_edges.clear();
  ix->live_slots = 0;
  for (auto& n : ix->nodes) n.in_slots.clear();
  for (uint32_t p = 0; p < ix->nodes.size(); ++p)
    for (auto& kv : ix->nodes[p].children)
      if (int rc = insert_edge(ix, p, kv.first, kv.second)) return rc;
  ix->full_upload_needed = true;
  return 0;
}

int new_node(kvdp_index* ix, uint32_t* out) {
  if (!ix->free_nodes.empty()) {  // a reclaimed id: its generation kept counting, so none of its old edges match
    *out = ix->free_nodes.back();
    ix->free_nodes.pop_back();
    return 0;
  }
  if ((int64_t)ix->nodes.size() >= ix->max_nodes)
    return fail(KVDP_ERANGE, "prefix index: node capacity %lld exhausted (live nodes)", (long long)ix->max_nodes);
  ix->nodes.emplace_back();
  if (ix->words > 1) ix->node_dirty_flag.push_back(0);
  *out = (uint32_t)ix->nodes.size() - 1;
  return 0;
}

// One reference to node `n` went away (a parent dropped its children, or a lookup entry was removed / replaced). At
// zero the node is gone for good, like an Rc: its own children lose a reference each, and its id goes on the free list.
void release_node(kvdp_index* ix, uint32_t first) {
  std::vector<uint32_t> work{first};
  while (!work.empty()) {
    const uint32_t n = work.back();
    work.pop_back();
    HostNode& h = ix->nodes[n];
    if (n == 0 || h.refs == 0 || --h.refs != 0) continue;
    for (auto& kv : h.children) work.push_back(kv.second);
    h.children.clear();
    ++h.gen;  // edges out of this id (device side) stop matching, also after the id is reused
    for (int w = 0; w < ix->max_workers; ++w)
      if (h.has(w)) --ix->member_nodes[w];  // a stale membership left by remove_worker dies with the node
    for (int k = 0; k < kMaxWorkerWords; ++k) h.workers[k] = 0;
    h.in_slots.clear();  // the edges that pointed here belong to dropped generations of their parents
    h.recent_uses.clear();
    ix->free_nodes.push_back(n);
  }
}

// Node `n` lost its last worker: none of its descendants is held through it any more (indexer.rs:366-369).
void drop_children(kvdp_index* ix, uint32_t n) {
  HostNode& h = ix->nodes[n];
  std::vector<uint32_t> kids;
  kids.reserve(h.children.size());
  for (auto& kv : h.children) kids.push_back(kv.second);
  h.children.clear();
  ++h.gen;  // device side: every edge out of this node stops matching
  for (uint32_t c : kids) release_node(ix, c);
}

int ensure_stage(kvdp_index* ix, size_t n) {
  if (n <= ix->stage_cap) return 0;
  if (ix->d_idx) cudaFree(ix->d_idx);
  if (ix->d_recs) cudaFree(ix->d_recs);
  ix->d_idx = nullptr;
  ix->d_recs = nullptr;
  ix->stage_cap = 0;
  const size_t cap = std::max<size_t>(n * 2, 4096);
  KVDP_CUDA(cudaMalloc(&ix->d_idx, cap * sizeof(uint32_t)));
  KVDP_CUDA(cudaMalloc(&ix->d_recs, cap * sizeof(EdgeRec)));
  ix->stage_cap = cap;
  return 0;
}

}  // namespace

extern "C" {

int kvdp_index_create(kvdp_index** out, int64_t max_nodes, int32_t max_workers, int32_t device) {
  if (!out) return fail(KVDP_EINVAL, "kvdp_index_create: NULL out");
  *out = nullptr;
  if (max_nodes < 2 || max_nodes > ((int64_t)1 << 31)) return fail(KVDP_EINVAL, "prefix index: max_nodes out of range");
  if (max_workers < 1 || max_workers > 64 * kMaxWorkerWords)
    return fail(KVDP_EINVAL, "prefix index: max_workers must be 1..%d", 64 * kMaxWorkerWords);
  kvdp_index* ix = new kvdp_index();
  ix->device = device;
  if (ix->device < 0) cudaGetDevice(&ix->device);
  ix->max_workers = max_workers;
  ix->words = (max_workers + 63) / 64;
  ix->lookup.resize(max_workers);
  ix->lookup_present.assign(max_workers, 0);
  ix->member_nodes.assign(max_workers, 0);
  ix->max_nodes = max_nodes;
  uint64_t cap = 1024;
  while (cap < (uint64_t)max_nodes * 4) cap <<= 1;  // load factor <= 0.25 for live edges, room for dead generations
  ix->capacity = cap;
  ix->h_edges.assign(cap, EdgeRec{});
  ix->edge_dirty_flag.assign(cap, 0);
  ix->nodes.reserve((size_t)std::min<int64_t>(max_nodes, 1 << 20));
  ix->nodes.emplace_back();  // root
  ix->nodes[0].refs = 1;     // never released
  if (ix->words > 1) ix->node_dirty_flag.push_back(0);
  DeviceGuard g(ix->device);
  cudaError_t e = cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMalloc(&ix->d_edges, sizeof(EdgeRec) * cap);
  if (e == cudaSuccess) e = cudaMemset(ix->d_edges, 0, sizeof(EdgeRec) * cap);
  if (e == cudaSuccess && ix->words > 1) {
    const size_t bytes = sizeof(uint64_t) * (size_t)max_nodes * (size_t)ix->words;
    e = cudaMalloc(&ix->d_wsets, bytes);
    if (e == cudaSuccess) e = cudaMemset(ix->d_wsets, 0, bytes);
  }
  if (e != cudaSuccess) {
    kvdp_index_destroy(ix);
    return fail((int)e, "prefix index: device allocation failed: %s", cudaGetErrorString(e));
  }
  ix->full_upload_needed = false;  // both sides are all-zero
  *out = ix;
  return 0;
}

int kvdp_index_destroy(kvdp_index* ix) {
  if (!ix) return 0;
  event_sink_forget_index(ix);  // an attached event publisher must not keep writing into freed memory
  DeviceGuard g(ix->device);
  if (ix->stream) {
    cudaStreamSynchronize(ix->stream);
    cudaStreamDestroy(ix->stream);
  }
  if (ix->d_edges) cudaFree(ix->d_edges);
  if (ix->d_wsets) cudaFree(ix->d_wsets);
  if (ix->d_idx) cudaFree(ix->d_idx);
  if (ix->d_recs) cudaFree(ix->d_recs);
  if (ix->d_io) cudaFree(ix->d_io);
  if (ix->h_io) cudaFreeHost(ix->h_io);
  delete ix;
  return 0;
}

kvdp_stream_t kvdp_index_stream(kvdp_index* ix) { return ix ? (kvdp_stream_t)ix->stream : nullptr; }

// indexer.rs:296-343. Returns 1 (and changes nothing) when the parent block is unknown for this worker.
int kvdp_index_apply_stored(kvdp_index* ix, int32_t worker, int32_t has_parent, uint64_t parent_hash,
                            const uint64_t* block_hashes, const uint64_t* tokens_hashes, int32_t n) {
  if (!ix) return fail(KVDP_EINVAL, "kvdp_index_apply_stored: NULL handle");
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  if (worker < 0 || worker >= ix->max_workers) return fail(KVDP_EINVAL, "prefix index: worker slot %d out of range", worker);
  if (n < 0 || (n > 0 && (!block_hashes || !tokens_hashes))) return fail(KVDP_EINVAL, "prefix index: bad block arrays");
  ix->lookup_present[worker] = true;
  auto& lk = ix->lookup[worker];
  uint32_t current = 0;
  if (has_parent) {
    auto it = lk.find(parent_hash);
    if (it == lk.end()) return 1;  // "Failed to find parent block; skipping store operation"
    current = it->second;
  }
  if (ix->live_slots + (uint64_t)n > ix->capacity / 2)
    if (int rc = rebuild_edges(ix)) return rc;
  for (int32_t i = 0; i < n; ++i) {
    uint32_t block;
    auto& children = ix->nodes[current].children;
    auto c = children.find(tokens_hashes[i]);
    if (c != children.end()) {
      block = c->second;
    } else {
      auto known = lk.find(block_hashes[i]);
      if (known != lk.end()) {
        block = known->second;  // the worker already has this block under another parent: share it
      } else if (int rc = new_node(ix, &block)) {
        return rc;
      }
      ix->nodes[current].children.emplace(tokens_hashes[i], block);
      ++ix->nodes[block].refs;
      if (int rc = insert_edge(ix, current, tokens_hashes[i], block)) return rc;
    }
    HostNode& b = ix->nodes[block];
    if (!b.has(worker)) {
      b.add(worker);
      ++ix->member_nodes[worker];
      mark_node(ix, block);
    }
    auto prev = lk.find(block_hashes[i]);
    if (prev == lk.end()) {
      lk.emplace(block_hashes[i], block);
      ++ix->nodes[block].refs;
    } else if (prev->second != block) {  // the engine reuses a block id for other content: the old node loses this entry
      const uint32_t old = prev->second;
      prev->second = block;
      ++ix->nodes[block].refs;
      release_node(ix, old);
    }
    current = block;
  }
  return 0;
}

// indexer.rs:344-376
int kvdp_index_apply_removed(kvdp_index* ix, int32_t worker, const uint64_t* block_hashes, int32_t n) {
  if (!ix) return fail(KVDP_EINVAL, "kvdp_index_apply_removed: NULL handle");
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  if (worker < 0 || worker >= ix->max_workers) return fail(KVDP_EINVAL, "prefix index: worker slot %d out of range", worker);
  if (n < 0 || (n > 0 && !block_hashes)) return fail(KVDP_EINVAL, "prefix index: bad block array");
  ix->lookup_present[worker] = true;
  auto& lk = ix->lookup[worker];
  for (int32_t i = 0; i < n; ++i) {
    auto it = lk.find(block_hashes[i]);
    if (it == lk.end()) continue;  // "Failed to find block to remove; skipping remove operation"
    const uint32_t n = it->second;
    {
      HostNode& b = ix->nodes[n];
      if (b.has(worker)) --ix->member_nodes[worker];
      b.drop(worker);
    }
    if (ix->nodes[n].none()) drop_children(ix, n);  // no worker holds this block, so none holds its descendants through it
    mark_node(ix, n);
    lk.erase(it);
    release_node(ix, n);
  }
  return 0;
}

// indexer.rs:380-386
int kvdp_index_remove_worker(kvdp_index* ix, int32_t worker) {
  if (!ix) return fail(KVDP_EINVAL, "kvdp_index_remove_worker: NULL handle");
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  if (worker < 0 || worker >= ix->max_workers) return fail(KVDP_EINVAL, "prefix index: worker slot %d out of range", worker);
  if (!ix->lookup_present[worker]) return 0;
  std::vector<uint32_t> held;
  held.reserve(ix->lookup[worker].size());
  for (auto& kv : ix->lookup[worker]) {
    HostNode& b = ix->nodes[kv.second];
    if (b.has(worker)) {
      b.drop(worker);
      --ix->member_nodes[worker];
      mark_node(ix, kv.second);
    }
    held.push_back(kv.second);
  }
  ix->lookup[worker].clear();
  ix->lookup_present[worker] = false;
  for (uint32_t n : held) release_node(ix, n);  // the table's references are gone (the blocks stay in the tree)
  return 0;
}

// Push every pending change to the device (asynchronous on the index's stream).
int kvdp_index_flush(kvdp_index* ix) {
  if (!ix) return fail(KVDP_EINVAL, "kvdp_index_flush: NULL handle");
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  DeviceGuard g(ix->device);
  if (ix->full_upload_needed) {
    KVDP_CUDA(cudaMemcpyAsync(ix->d_edges, ix->h_edges.data(), sizeof(EdgeRec) * ix->capacity, cudaMemcpyHostToDevice,
                              ix->stream));
    KVDP_CUDA(cudaStreamSynchronize(ix->stream));  // pageable source: keep the host table stable until copied
    for (uint32_t s : ix->dirty_edges) ix->edge_dirty_flag[s] = 0;
    ix->dirty_edges.clear();
    ix->full_upload_needed = false;
  }
  if (ix->words > 1 && !ix->dirty_nodes.empty()) {
    const size_t nn = ix->dirty_nodes.size();
    if (int rc = ensure_stage(ix, nn * (size_t)ix->words)) return rc;  // d_recs holds 32 bytes per staged entry
    std::vector<uint64_t> sets(nn * (size_t)ix->words);
    for (size_t i = 0; i < nn; ++i)
      for (int k = 0; k < ix->words; ++k) sets[i * ix->words + k] = ix->nodes[ix->dirty_nodes[i]].workers[k];
    KVDP_CUDA(cudaMemcpyAsync(ix->d_idx, ix->dirty_nodes.data(), nn * sizeof(uint32_t), cudaMemcpyHostToDevice, ix->stream));
    KVDP_CUDA(cudaMemcpyAsync(ix->d_recs, sets.data(), sets.size() * sizeof(uint64_t), cudaMemcpyHostToDevice, ix->stream));
    kvdp_index_scatter_words_kernel<<<(unsigned)((sets.size() + 255) / 256), 256, 0, ix->stream>>>(
        ix->d_wsets, ix->d_idx, (const uint64_t*)ix->d_recs, (int)nn, ix->words);
    KVDP_CUDA(cudaStreamSynchronize(ix->stream));
    for (uint32_t n : ix->dirty_nodes) ix->node_dirty_flag[n] = 0;
    ix->dirty_nodes.clear();
    KVDP_CUDA(cudaGetLastError());
  }
  const size_t ne = ix->dirty_edges.size();
  if (ne == 0) return 0;
  if (int rc = ensure_stage(ix, ne)) return rc;
  std::vector<EdgeRec> recs(ne);
  for (size_t i = 0; i < ne; ++i) recs[i] = ix->h_edges[ix->dirty_edges[i]];
  KVDP_CUDA(cudaMemcpyAsync(ix->d_idx, ix->dirty_edges.data(), ne * sizeof(uint32_t), cudaMemcpyHostToDevice, ix->stream));
  KVDP_CUDA(cudaMemcpyAsync(ix->d_recs, recs.data(), ne * sizeof(EdgeRec), cudaMemcpyHostToDevice, ix->stream));
  kvdp_index_scatter_kernel<EdgeRec><<<(unsigned)((ne + 255) / 256), 256, 0, ix->stream>>>(
      ix->d_edges, ix->d_idx, (const EdgeRec*)ix->d_recs, (int)ne);
  KVDP_CUDA(cudaStreamSynchronize(ix->stream));
  for (uint32_t s : ix->dirty_edges) ix->edge_dirty_flag[s] = 0;
  ix->dirty_edges.clear();
  KVDP_CUDA(cudaGetLastError());
  return 0;
}

// Batched RadixTree::find_matches (indexer.rs:239-277) for n_req requests given as CSR sequences of LocalBlockHash on
// the device. scores is [n_req][num_workers] (u32, worker slot order); matched_depth[n_req] may be NULL. Pending
// changes are flushed first. Asynchronous on `stream` (NULL = the index's own stream).
int kvdp_index_find_matches(kvdp_index* ix, const uint64_t* d_hashes, const int64_t* d_seq_offsets, int64_t n_req,
                            int32_t early_exit, int32_t num_workers, uint32_t* d_scores, int32_t* d_matched_depth,
                            kvdp_stream_t stream_) {
  if (!ix) return fail(KVDP_EINVAL, "kvdp_index_find_matches: NULL handle");
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  if (n_req < 0) return fail(KVDP_EINVAL, "prefix index: negative n_req");
  if (num_workers < 1 || num_workers > ix->max_workers) return fail(KVDP_EINVAL, "prefix index: num_workers out of range");
  if (n_req == 0) return 0;
  if (!d_seq_offsets || !d_scores) return fail(KVDP_EINVAL, "prefix index: NULL pointer");
  if (int rc = kvdp_index_flush(ix)) return rc;
  DeviceGuard g(ix->device);
  cudaStream_t stream = stream_ ? (cudaStream_t)stream_ : ix->stream;
  const dim3 grid((unsigned)((n_req + 127) / 128), (unsigned)(ix->words > 1 ? (num_workers + 63) / 64 : 1));
  if (ix->words > 1)
    kvdp_index_find_kernel<true><<<grid, 128, 0, stream>>>(ix->d_edges, ix->capacity - 1, ix->nodes[0].gen, d_hashes,
                                                           d_seq_offsets, n_req, early_exit, num_workers, d_scores,
                                                           d_matched_depth, ix->d_wsets, ix->words);
  else
    kvdp_index_find_kernel<false><<<grid, 128, 0, stream>>>(ix->d_edges, ix->capacity - 1, ix->nodes[0].gen, d_hashes,
                                                            d_seq_offsets, n_req, early_exit, num_workers, d_scores,
                                                            d_matched_depth, nullptr, 1);
  KVDP_CUDA(cudaGetLastError());
  return 0;
}

// Host-buffer form: hashes, offsets in; scores (and depths) out; synchronises. One pinned staging buffer, ONE upload and
// ONE download per call instead of a copy per array, so a single-request lookup costs little more than the launch.
int kvdp_index_find_matches_host(kvdp_index* ix, const uint64_t* h_hashes, const int64_t* h_seq_offsets, int64_t n_req,
                                 int32_t early_exit, int32_t num_workers, uint32_t* h_scores, int32_t* h_matched_depth) {
  if (!ix) return fail(KVDP_EINVAL, "kvdp_index_find_matches_host: NULL handle");
  if (n_req < 0) return fail(KVDP_EINVAL, "prefix index: negative n_req");
  if (num_workers < 1 || num_workers > ix->max_workers) return fail(KVDP_EINVAL, "prefix index: num_workers out of range");
  if (n_req == 0) return 0;
  if (!h_seq_offsets || !h_scores) return fail(KVDP_EINVAL, "prefix index: NULL pointer");
  const int64_t n_hash = h_seq_offsets[n_req] - h_seq_offsets[0];
  if (h_seq_offsets[0] != 0 || n_hash < 0 || (n_hash > 0 && !h_hashes))
    return fail(KVDP_EINVAL, "prefix index: offsets must start at 0 and hashes must be given");
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  DeviceGuard g(ix->device);
  const size_t in_hash = (size_t)n_hash * 8, in_off = (size_t)(n_req + 1) * 8;
  const size_t out_scores = (size_t)n_req * num_workers * 4, out_depth = (size_t)n_req * 4;
  const size_t in_bytes = (in_hash + in_off + 15) & ~(size_t)15;
  const size_t need = in_bytes + out_scores + out_depth;
  if (need > ix->io_cap) {
    KVDP_CUDA(cudaStreamSynchronize(ix->stream));
    if (ix->d_io) cudaFree(ix->d_io);
    if (ix->h_io) cudaFreeHost(ix->h_io);
    ix->d_io = ix->h_io = nullptr;
    ix->io_cap = 0;
    const size_t cap = std::max<size_t>(need * 2, 1 << 16);
    KVDP_CUDA(cudaMalloc((void**)&ix->d_io, cap));
    KVDP_CUDA(cudaHostAlloc((void**)&ix->h_io, cap, cudaHostAllocDefault));
    ix->io_cap = cap;
  }
  memcpy(ix->h_io, h_seq_offsets, in_off);          // offsets first: 8-byte aligned either way
  if (in_hash) memcpy(ix->h_io + in_off, h_hashes, in_hash);
  KVDP_CUDA(cudaMemcpyAsync(ix->d_io, ix->h_io, in_off + in_hash, cudaMemcpyHostToDevice, ix->stream));
  uint32_t* d_scores = (uint32_t*)(ix->d_io + in_bytes);
  int32_t* d_depth = (int32_t*)(ix->d_io + in_bytes + out_scores);
  if (int rc = kvdp_index_find_matches(ix, (const uint64_t*)(ix->d_io + in_off), (const int64_t*)ix->d_io, n_req,
                                       early_exit, num_workers, d_scores, d_depth, nullptr))
    return rc;
  KVDP_CUDA(cudaMemcpyAsync(ix->h_io + in_bytes, d_scores, out_scores + out_depth, cudaMemcpyDeviceToHost, ix->stream));
  KVDP_CUDA(cudaStreamSynchronize(ix->stream));
  memcpy(h_scores, ix->h_io + in_bytes, out_scores);
  if (h_matched_depth) memcpy(h_matched_depth, ix->h_io + in_bytes + out_scores, out_depth);
  return 0;
}

// KvIndexer::find_matches_for_request (indexer.rs:639-651) for a batch of token sequences held by the host: hash the full
// blocks (seed 1337, un-chained) and walk the tree in the same round trip — one upload, scan + hash + find kernels on
// the index's stream, one download.
int kvdp_index_find_matches_for_tokens_host(kvdp_index* ix, const uint32_t* h_tokens, const int64_t* h_offsets,
                                            int64_t n_req, int32_t block_size, int32_t early_exit, int32_t num_workers,
                                            uint32_t* h_scores, int32_t* h_matched_depth) {
  if (!ix) return fail(KVDP_EINVAL, "kvdp_index_find_matches_for_tokens_host: NULL handle");
  if (n_req < 0 || block_size < 1) return fail(KVDP_EINVAL, "prefix index: bad n_req / block_size");
  if (num_workers < 1 || num_workers > ix->max_workers) return fail(KVDP_EINVAL, "prefix index: num_workers out of range");
  if (n_req == 0) return 0;
  if (!h_offsets || !h_scores) return fail(KVDP_EINVAL, "prefix index: NULL pointer");
  const int64_t total = h_offsets[n_req];
  if (h_offsets[0] != 0 || total < 0 || (total > 0 && !h_tokens))
    return fail(KVDP_EINVAL, "prefix index: offsets must start at 0 and tokens must be given");
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  DeviceGuard g(ix->device);
  auto up16 = [](size_t v) { return (v + 15) & ~(size_t)15; };
  const size_t b_off = (size_t)(n_req + 1) * 8, b_tok = up16((size_t)total * 4);
  const int64_t max_blocks = total / block_size;
  const size_t b_bh = up16((size_t)std::max<int64_t>(max_blocks, 1) * 8), b_ws = up16(kvdp_hash_workspace_bytes(n_req));
  const size_t out_scores = (size_t)n_req * num_workers * 4, out_depth = (size_t)n_req * 4;
  const size_t o_tok = up16(b_off), o_boff = o_tok + b_tok, o_bh = o_boff + up16(b_off), o_ws = o_bh + b_bh;
  const size_t o_out = (o_ws + b_ws + 255) & ~(size_t)255;
  const size_t need = o_out + out_scores + out_depth;
  if (need > ix->io_cap) {
    KVDP_CUDA(cudaStreamSynchronize(ix->stream));
    if (ix->d_io) cudaFree(ix->d_io);
    if (ix->h_io) cudaFreeHost(ix->h_io);
    ix->d_io = ix->h_io = nullptr;
    ix->io_cap = 0;
    const size_t cap = std::max<size_t>(need * 2, 1 << 16);
    KVDP_CUDA(cudaMalloc((void**)&ix->d_io, cap));
    KVDP_CUDA(cudaHostAlloc((void**)&ix->h_io, cap, cudaHostAllocDefault));
    ix->io_cap = cap;
  }
  memcpy(ix->h_io, h_offsets, b_off);
  if (total) memcpy(ix->h_io + o_tok, h_tokens, (size_t)total * 4);
  KVDP_CUDA(cudaMemcpyAsync(ix->d_io, ix->h_io, o_tok + (size_t)total * 4, cudaMemcpyHostToDevice, ix->stream));
  int64_t* d_boff = (int64_t*)(ix->d_io + o_boff);
  uint64_t* d_bh = (uint64_t*)(ix->d_io + o_bh);
  if (int rc = kvdp_hash_token_blocks((const uint32_t*)(ix->d_io + o_tok), (const int64_t*)ix->d_io, n_req, total, block_size,
                                      1337, d_boff, d_bh, nullptr, ix->d_io + o_ws, b_ws, 0, (kvdp_stream_t)ix->stream))
    return rc;
  uint32_t* d_scores = (uint32_t*)(ix->d_io + o_out);
  int32_t* d_depth = (int32_t*)(ix->d_io + o_out + out_scores);
  if (int rc = kvdp_index_find_matches(ix, d_bh, d_boff, n_req, early_exit, num_workers, d_scores, d_depth, nullptr))
    return rc;
  KVDP_CUDA(cudaMemcpyAsync(ix->h_io + o_out, d_scores, out_scores + out_depth, cudaMemcpyDeviceToHost, ix->stream));
  KVDP_CUDA(cudaStreamSynchronize(ix->stream));
  memcpy(h_scores, ix->h_io + o_out, out_scores);
  if (h_matched_depth) memcpy(h_matched_depth, ix->h_io + o_out + out_scores, out_depth);
  return 0;
}

// Access-frequency bookkeeping of RadixTree::find_matches when the tree was built with an expiration duration
// (indexer.rs:217-223, 251-267): every block on the matched path keeps the times it was visited; a visit first drops
// the entries older than `expiration_ns`, reports how many are left (zeros are not reported, :429-436) and then records
// itself. This is per-request, order-dependent host state, so it lives with the host tree; the scores still come from
// the GPU walk. Returns the number of frequencies written (they are non-increasing along the path), or a negative
// KVDP_E* code; `cap` must hold one entry per block of the sequence.
int64_t kvdp_index_touch(kvdp_index* ix, const uint64_t* h_hashes, int64_t n, int32_t early_exit, int64_t now_ns,
                         int64_t expiration_ns, uint64_t* h_frequencies, int64_t cap) {
  if (!ix) return fail(KVDP_EINVAL, "kvdp_index_touch: NULL handle");
  if (n < 0 || (n > 0 && !h_hashes) || cap < n || (n > 0 && !h_frequencies))
    return fail(KVDP_EINVAL, "kvdp_index_touch: bad arguments (cap must be >= n)");
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  uint32_t current = 0;
  int64_t written = 0;
  for (int64_t i = 0; i < n; ++i) {
    auto it = ix->nodes[current].children.find(h_hashes[i]);
    if (it == ix->nodes[current].children.end()) break;
    HostNode& b = ix->nodes[it->second];
    while (!b.recent_uses.empty() && now_ns - b.recent_uses.front() > expiration_ns) b.recent_uses.pop_front();
    if (!b.recent_uses.empty()) h_frequencies[written++] = (uint64_t)b.recent_uses.size();
    b.recent_uses.push_back(now_ns);
    if (early_exit && b.count() == 1) break;
    current = it->second;
  }
  return written;
}

// sizes of the host-side structures, for the reference tests' assertions:
// what = 0: number of nodes; 1: live edges out of `arg` (node id); 2: entries in worker `arg`'s lookup table;
//        3: popcount of node `arg`'s worker set; 4: child of the root for tokens hash `arg` (node id or -1);
//        5: workers that ever had a lookup table and still do; 6: nodes whose worker set holds slot `arg`
int64_t kvdp_index_stat(kvdp_index* ix, int32_t what, uint64_t arg) {
  if (!ix) return -1;
  std::lock_guard<std::recursive_mutex> guard(ix->mu);
  switch (what) {
    case 0: return (int64_t)(ix->nodes.size() - ix->free_nodes.size());
    case 1: return arg < ix->nodes.size() ? (int64_t)ix->nodes[arg].children.size() : -1;
    case 2: return arg < (uint64_t)ix->max_workers ? (int64_t)ix->lookup[arg].size() : -1;
    case 3: return arg < ix->nodes.size() ? (int64_t)ix->nodes[arg].count() : -1;
    case 4: {
      auto it = ix->nodes[0].children.find(arg);
      return it == ix->nodes[0].children.end() ? -1 : (int64_t)it->second;
    }
    case 5: {
      int64_t c = 0;
      for (int w = 0; w < ix->max_workers; ++w) c += ix->lookup_present[w] ? 1 : 0;
      return c;
    }
    case 6: return arg < (uint64_t)ix->max_workers ? ix->member_nodes[arg] : -1;
    case 7: return (int64_t)ix->free_nodes.size();
    default: return -1;
  }
}

}  // extern "C"
