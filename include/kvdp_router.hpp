// kvdp_router.hpp — C++17 host layer for the rows either side of the block path, over the C ABI of kvdp.h:
//
//   RadixTree, KvIndexer, KvIndexerSharded   lib/llm/src/kv_router/indexer.rs:187-387, 498-667, 677-884
//   KvEventPublisher                          lib/llm/src/kv_router/publisher.rs:34-58 + lib/bindings/c/src/lib.rs:50-311
//   KvBlock, AvailableBlocks, ReservedBlocks, KvStorageManager
//                                             lib/llm/src/kv.rs:47-91, kv/reuse.rs:40-660, kv/reserved.rs:15-160,
//                                             kv/manager.rs:15-142
//
// Same names and call semantics as the Rust types; the tokio actors become direct synchronous calls and Rust's
// drop-returns-to-pool becomes shared_ptr deleters. The reference's hashing and tree walks run on the GPU through
// libkvdp.so (batched entry points are added next to the per-request ones); the pool is plain host logic, as it is there.
// Header-only; link with -lkvdp.
#pragma once
#include <chrono>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <set>
#include <unordered_map>

#include "kvdp.hpp"

namespace kvdp {

using WorkerId = int64_t;

// ------------------------------------------------------------------------------------------------ router events
struct KvCacheStoredBlockData {  // kv_router/protocols.rs:107-113
  uint64_t block_hash;           // ExternalSequenceBlockHash: the engine's id for the block
  uint64_t tokens_hash;          // LocalBlockHash: seed-1337 hash of the block's tokens
};
struct KvCacheStoreData {        // protocols.rs:98-104
  std::optional<uint64_t> parent_hash;
  std::vector<KvCacheStoredBlockData> blocks;
};
struct KvCacheRemoveData {       // protocols.rs:116-120
  std::vector<uint64_t> block_hashes;
};
struct KvCacheEvent {            // protocols.rs:79-95: exactly one of stored / removed is set
  uint64_t event_id = 0;
  std::optional<KvCacheStoreData> stored;
  std::optional<KvCacheRemoveData> removed;
};
struct RouterEvent {             // indexer.rs:139-160
  WorkerId worker_id = 0;
  KvCacheEvent event;
};
struct OverlapScores {           // indexer.rs:390-436 (frequencies: only with an expiration duration)
  std::map<WorkerId, uint32_t> scores;
  std::vector<size_t> frequencies;
};

// ------------------------------------------------------------------------------------------------ prefix index
class RadixTree {  // indexer.rs:187-387
 public:
  // `expiration_s`: RadixTree::new_with_frequency (indexer.rs:217-223); lookups then also report visit frequencies
  // `max_workers` (<= 512): workers the index holds at a time (the reference's set is unbounded, indexer.rs:163-170)
  explicit RadixTree(int64_t max_nodes = 1 << 20, int device = -1, std::optional<double> expiration_s = std::nullopt,
                     int max_workers = 64)
      : expiration_s_(expiration_s), max_workers_(max_workers) {
    if (kvdp_index_create(&ix_, max_nodes, max_workers_, device) != 0) raise(kvdp_last_error());
    slot_worker_.assign(max_workers_, std::nullopt);
  }
  ~RadixTree() { kvdp_index_destroy(ix_); }
  RadixTree(const RadixTree&) = delete;
  RadixTree& operator=(const RadixTree&) = delete;

  void apply_event(const RouterEvent& e) {  // indexer.rs:284-378
    const int slot = slot_of(e.worker_id, true);
    if (e.event.stored) {
      const auto& s = *e.event.stored;
      std::vector<uint64_t> bh, th;
      for (const auto& b : s.blocks) {
        bh.push_back(b.block_hash);
        th.push_back(b.tokens_hash);
      }
      // rc == 1: parent unknown for this worker; the reference logs a warning and skips the event (:305-315)
      if (kvdp_index_apply_stored(ix_, slot, s.parent_hash ? 1 : 0, s.parent_hash.value_or(0), bh.data(), th.data(),
                                  (int32_t)bh.size()) < 0)
        raise(kvdp_last_error());
    } else if (e.event.removed) {
      const auto& r = e.event.removed->block_hashes;
      if (kvdp_index_apply_removed(ix_, slot, r.data(), (int32_t)r.size()) != 0) raise(kvdp_last_error());
    }
  }
  void remove_worker(WorkerId w) {  // indexer.rs:380-386
    const int slot = slot_of(w, false);
    if (slot < 0) return;
    if (kvdp_index_remove_worker(ix_, slot) != 0) raise(kvdp_last_error());
    if (kvdp_index_stat(ix_, 6, (uint64_t)slot) == 0) {  // nothing stale left under this worker: the slot is free again
      worker_slot_.erase(w);
      slot_worker_[slot].reset();
    }
  }
  // indexer.rs:239-277. `now_s`: clock override for tests (default: steady clock)
  OverlapScores find_matches(const std::vector<uint64_t>& sequence, bool early_exit = false,
                             std::optional<double> now_s = std::nullopt) {
    OverlapScores out = find_matches_batch({sequence}, early_exit)[0];
    if (expiration_s_) {
      const int64_t now_ns =
          now_s ? (int64_t)(*now_s * 1e9 + 0.5)
                : (int64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(
                      std::chrono::steady_clock::now().time_since_epoch()).count();
      std::vector<uint64_t> f(std::max<size_t>(sequence.size(), 1));
      const int64_t n = kvdp_index_touch(ix_, sequence.data(), (int64_t)sequence.size(), early_exit ? 1 : 0, now_ns,
                                         (int64_t)(*expiration_s_ * 1e9 + 0.5), f.data(), (int64_t)f.size());
      if (n < 0) raise(kvdp_last_error());
      out.frequencies.assign(f.begin(), f.begin() + n);
    }
    return out;
  }
  // Many requests in one launch (new).
  std::vector<OverlapScores> find_matches_batch(const std::vector<std::vector<uint64_t>>& sequences,
                                                bool early_exit = false) {
    const int64_t n = (int64_t)sequences.size();
    std::vector<OverlapScores> out((size_t)n);
    if (n == 0) return out;
    std::vector<int64_t> offs((size_t)n + 1, 0);
    for (int64_t i = 0; i < n; ++i) offs[i + 1] = offs[i] + (int64_t)sequences[i].size();
    std::vector<uint64_t> flat((size_t)std::max<int64_t>(offs[n], 1), 0);
    for (int64_t i = 0; i < n; ++i) std::copy(sequences[i].begin(), sequences[i].end(), flat.begin() + offs[i]);
    const int nw = workers_in_use();
    std::vector<uint32_t> scores((size_t)n * nw);
    if (kvdp_index_find_matches_host(ix_, flat.data(), offs.data(), n, early_exit ? 1 : 0, nw, scores.data(), nullptr) != 0)
      raise(kvdp_last_error());
    for (int64_t i = 0; i < n; ++i)
      for (int s = 0; s < nw; ++s)
        if (scores[(size_t)i * nw + s] && slot_worker_[s]) out[i].scores[*slot_worker_[s]] = scores[(size_t)i * nw + s];
    return out;
  }
  // Token sequences in: seed-1337 block hashing and the walk in one round trip to the GPU (new).
  std::vector<OverlapScores> find_matches_for_tokens(const std::vector<std::vector<Token>>& requests, size_t block_size) {
    const int64_t n = (int64_t)requests.size();
    std::vector<OverlapScores> out((size_t)n);
    if (n == 0) return out;
    std::vector<int64_t> offs((size_t)n + 1, 0);
    for (int64_t i = 0; i < n; ++i) offs[i + 1] = offs[i] + (int64_t)requests[i].size();
    std::vector<Token> flat((size_t)std::max<int64_t>(offs[n], 1), 0);
    for (int64_t i = 0; i < n; ++i) std::copy(requests[i].begin(), requests[i].end(), flat.begin() + offs[i]);
    const int nw = workers_in_use();
    std::vector<uint32_t> scores((size_t)n * nw);
    if (kvdp_index_find_matches_for_tokens_host(ix_, flat.data(), offs.data(), n, (int32_t)block_size, 0, nw, scores.data(),
                                                nullptr) != 0)
      raise(kvdp_last_error());
    for (int64_t i = 0; i < n; ++i)
      for (int s = 0; s < nw; ++s)
        if (scores[(size_t)i * nw + s] && slot_worker_[s]) out[i].scores[*slot_worker_[s]] = scores[(size_t)i * nw + s];
    return out;
  }
  int64_t stat(int what, uint64_t arg = 0) const { return kvdp_index_stat(ix_, what, arg); }
  kvdp_index* handle() const { return ix_; }
  int slot_of(WorkerId w, bool create) {
    auto it = worker_slot_.find(w);
    if (it != worker_slot_.end()) return it->second;
    if (!create) return -1;
    for (int s = 0; s < max_workers_; ++s)
      if (!slot_worker_[s]) {
        slot_worker_[s] = w;
        worker_slot_[w] = s;
        return s;
      }
    raise("the prefix index is full: it was created for " + std::to_string(max_workers_) + " workers at a time");
    return -1;
  }

 private:
  int workers_in_use() const {  // highest bound slot + 1 (at least 1): the width of the score rows
    int nw = 1;
    for (size_t s = 0; s < slot_worker_.size(); ++s)
      if (slot_worker_[s]) nw = (int)s + 1;
    return nw;
  }
  kvdp_index* ix_ = nullptr;
  std::optional<double> expiration_s_;
  int max_workers_ = 64;
  std::unordered_map<WorkerId, int> worker_slot_;
  std::vector<std::optional<WorkerId>> slot_worker_;
};

class KvIndexer {  // indexer.rs:498-667, without the actor
 public:
  explicit KvIndexer(size_t kv_block_size, int64_t max_nodes = 1 << 20, int device = -1)
      : kv_block_size_(kv_block_size), tree_(max_nodes, device) {}
  size_t block_size() const { return kv_block_size_; }
  void apply_event(const RouterEvent& e) { tree_.apply_event(e); }
  void remove_worker(WorkerId w) { tree_.remove_worker(w); }
  OverlapScores find_matches(const std::vector<uint64_t>& sequence) { return tree_.find_matches(sequence, false); }
  OverlapScores find_matches_for_request(const std::vector<Token>& tokens) {  // indexer.rs:639-651
    return tree_.find_matches_for_tokens({tokens}, kv_block_size_)[0];
  }
  RadixTree& tree() { return tree_; }

 private:
  size_t kv_block_size_;
  RadixTree tree_;
};

class KvIndexerSharded {  // indexer.rs:677-884: a worker goes to the shard that has the fewest workers at its first event
 public:
  KvIndexerSharded(size_t num_shards, size_t kv_block_size, int64_t max_nodes = 1 << 20, int device = -1)
      : kv_block_size_(kv_block_size), counts_(num_shards, 0) {
    if (num_shards < 1) raise("num_shards must be at least 1");
    for (size_t i = 0; i < num_shards; ++i) shards_.push_back(std::make_unique<RadixTree>(max_nodes, device));
  }
  size_t block_size() const { return kv_block_size_; }
  void apply_event(const RouterEvent& e) {
    auto it = assignment_.find(e.worker_id);
    if (it == assignment_.end()) {
      size_t best = 0;
      for (size_t i = 1; i < counts_.size(); ++i)
        if (counts_[i] < counts_[best]) best = i;
      it = assignment_.emplace(e.worker_id, best).first;
      ++counts_[best];
    }
    shards_[it->second]->apply_event(e);
  }
  void remove_worker(WorkerId w) {
    auto it = assignment_.find(w);
    if (it == assignment_.end()) return;
    const size_t shard = it->second;
    assignment_.erase(it);
    --counts_[shard];
    shards_[shard]->remove_worker(w);
  }
  OverlapScores find_matches(const std::vector<uint64_t>& sequence) {
    OverlapScores out;
    for (auto& s : shards_) {
      OverlapScores part = s->find_matches(sequence, false);
      for (auto& kv : part.scores) out.scores[kv.first] = kv.second;
    }
    return out;
  }
  OverlapScores find_matches_for_request(const std::vector<Token>& tokens) {
    OverlapScores out;
    for (auto& s : shards_)
      for (auto& kv : s->find_matches_for_tokens({tokens}, kv_block_size_)[0].scores) out.scores[kv.first] = kv.second;
    return out;
  }
  const std::vector<size_t>& worker_counts() const { return counts_; }

 private:
  size_t kv_block_size_;
  std::vector<std::unique_ptr<RadixTree>> shards_;
  std::unordered_map<WorkerId, size_t> assignment_;
  std::vector<size_t> counts_;
};

// ------------------------------------------------------------------------------------------------ event publisher
// One per process, like the reference's KV_PUB OnceCell (lib/bindings/c/src/lib.rs:26-29).
class KvEventPublisher {
 public:
  KvEventPublisher(const std::string& ns, const std::string& component, WorkerId worker_id, size_t kv_block_size)
      : worker_id_(worker_id), kv_block_size_(kv_block_size) {
    if (dynamo_llm_init(ns.c_str(), component.c_str(), worker_id, (uint32_t)kv_block_size) != DYNAMO_LLM_OK)
      raise("dynamo_llm_init failed");
  }
  size_t kv_block_size() const { return kv_block_size_; }
  // kv.rs:151-178 / lib.rs:260-295: block i owns num_block_tokens[i] tokens of token_ids; a block that is not full ends the list
  void publish_stored(uint64_t event_id, const std::vector<Token>& token_ids, const std::vector<size_t>& num_block_tokens,
                      const std::vector<uint64_t>& block_hashes, uint64_t lora_id,
                      std::optional<uint64_t> parent_hash = std::nullopt) {
    const size_t n = std::min(num_block_tokens.size(), block_hashes.size());
    const uint64_t parent = parent_hash.value_or(0);
    if (dynamo_kv_event_publish_stored(event_id, token_ids.data(), num_block_tokens.data(), block_hashes.data(), n,
                                       parent_hash ? &parent : nullptr, lora_id) != DYNAMO_LLM_OK)
      raise("dynamo_kv_event_publish_stored failed");
  }
  void publish_removed(uint64_t event_id, const std::vector<uint64_t>& block_hashes) {
    if (dynamo_kv_event_publish_removed(event_id, block_hashes.data(), block_hashes.size()) != DYNAMO_LLM_OK)
      raise("dynamo_kv_event_publish_removed failed");
  }
  // sinks (no reference counterpart; the reference publishes on NATS)
  void attach_index(RadixTree* tree) {
    if (kvdp_event_sink_index(tree ? tree->handle() : nullptr, tree ? tree->slot_of(worker_id_, true) : 0) != 0)
      raise(kvdp_last_error());
  }
  void on_event(kvdp_event_cb cb, void* user) { kvdp_event_sink_callback(cb, user); }
  void flush() { kvdp_event_flush(); }
  std::optional<std::string> poll() {  // next RouterEvent JSON from the outbox
    size_t len = 0;
    std::string buf(4096, '\0');
    int rc = kvdp_event_poll(buf.data(), buf.size(), &len);
    if (rc == KVDP_ERANGE) {
      buf.assign(len, '\0');
      rc = kvdp_event_poll(buf.data(), buf.size(), &len);
    }
    if (rc < 0) raise(kvdp_last_error());
    if (rc == 0) return std::nullopt;
    buf.resize(len);
    return buf;
  }
  void shutdown() { dynamo_llm_shutdown(); }

 private:
  WorkerId worker_id_;
  size_t kv_block_size_;
};

// ------------------------------------------------------------------------------------------------ reuse pool
struct KvBlock {  // kv.rs:47-91
  TokenBlock token_block;  // default: sequence_hash 0 = un-initialised
  uint32_t priority = 0;
  uint64_t return_tick = 0;
  int64_t block_id = -1;  // slot in the KV cache slabs (new: what the copy kernels index with; kv.rs:54-63 leaves it open)
  void update_token_block(TokenBlock tb) { token_block = std::move(tb); }
  void reset() {
    token_block = TokenBlock{};
    priority = 0;
    return_tick = 0;
  }
};

struct UpdateBlock {  // reuse.rs:596-600
  SequenceHash hash;
  std::optional<uint32_t> priority;
};

class AvailableBlocks;
// A block on loan: when the last copy of the handle goes away the block returns to its pool (PoolItem / SharedPoolItem).
using PoolItem = std::shared_ptr<KvBlock>;

class AvailableBlocks : public std::enable_shared_from_this<AvailableBlocks> {  // reuse.rs:40-660
 public:
  static std::shared_ptr<AvailableBlocks> create() { return std::shared_ptr<AvailableBlocks>(new AvailableBlocks()); }
  uint64_t total_blocks() const { return total_; }
  uint64_t available_blocks() const { return available_; }
  bool is_active() const { return true; }
  void fence() {}  // the reference waits for its actor to drain; calls here are synchronous already

  void insert(KvBlock block) {  // handle_insert, reuse.rs:505-517
    ++available_;
    ++total_;
    block.return_tick = ++return_tick_;
    place(std::make_unique<KvBlock>(std::move(block)));
  }
  // Take the blocks whose hashes match, in order, stopping at the first miss (reuse.rs:355-374).
  std::vector<PoolItem> match_blocks(const std::vector<SequenceHash>& hashes) {
    std::vector<PoolItem> out;
    for (SequenceHash h : hashes) {
      auto b = take_with_hash(h);
      if (!b) break;
      out.push_back(lend(std::move(b)));
    }
    available_ -= out.size();
    return out;
  }
  std::vector<PoolItem> match_token_blocks(const std::vector<TokenBlock>& blocks) {
    std::vector<SequenceHash> h;
    for (const auto& b : blocks) h.push_back(b.sequence_hash);
    return match_blocks(h);
  }
  // Un-initialised blocks first, then lowest (priority, return tick) (reuse.rs:400-420); fewer than `count` if empty.
  std::vector<PoolItem> take_blocks(size_t count) {
    std::vector<PoolItem> out;
    for (size_t i = 0; i < count; ++i) {
      std::unique_ptr<KvBlock> b;
      if (!uninitialized_.empty()) {
        b = std::move(uninitialized_.front());
        uninitialized_.pop_front();
      } else if (!priority_set_.empty()) {
        const SequenceHash h = priority_set_.begin()->second;
        priority_set_.erase(priority_set_.begin());
        auto it = lookup_.find(h);
        if (it == lookup_.end()) raise("block from priority set not found in lookup map");
        b = std::move(it->second);
        lookup_.erase(it);
      } else {
        break;
      }
      out.push_back(lend(std::move(b)));
    }
    available_ -= out.size();
    return out;
  }
  void update_single(const UpdateBlock& u) { update_multiple({u}); }
  void update_multiple(const std::vector<UpdateBlock>& updates) {  // reuse.rs:537-552
    for (const auto& u : updates) {
      auto b = take_with_hash(u.hash);
      if (!b) continue;
      if (u.priority) b->priority = *u.priority;
      place(std::move(b));
    }
  }
  void reset(const std::vector<SequenceHash>& hashes) {  // reuse.rs:554-561
    for (SequenceHash h : hashes) {
      auto b = take_with_hash(h);
      if (!b) continue;
      b->reset();
      place(std::move(b));
    }
  }
  void reset_all() {  // reuse.rs:563-573
    while (!priority_set_.empty()) {
      const SequenceHash h = priority_set_.begin()->second;
      priority_set_.erase(priority_set_.begin());
      auto it = lookup_.find(h);
      if (it == lookup_.end()) raise("block from priority set not found in lookup map");
      auto b = std::move(it->second);
      lookup_.erase(it);
      b->reset();
      place(std::move(b));
    }
  }

 private:
  AvailableBlocks() = default;
  void place(std::unique_ptr<KvBlock> b) {  // one entry per hash; hash 0 and duplicates wait in the un-initialised queue
    const SequenceHash h = b->token_block.sequence_hash;
    if (h == 0 || lookup_.count(h)) {
      uninitialized_.push_back(std::move(b));
      return;
    }
    const auto key = std::make_pair(b->priority, b->return_tick);
    if (!priority_set_.emplace(key, h).second) raise("fatal error: multiple entries for the same key in priority set");
    lookup_.emplace(h, std::move(b));
  }
  std::unique_ptr<KvBlock> take_with_hash(SequenceHash h) {
    auto it = lookup_.find(h);
    if (it == lookup_.end()) return nullptr;
    auto b = std::move(it->second);
    lookup_.erase(it);
    priority_set_.erase(std::make_pair(b->priority, b->return_tick));
    return b;
  }
  PoolItem lend(std::unique_ptr<KvBlock> b) {
    std::weak_ptr<AvailableBlocks> pool = weak_from_this();
    return PoolItem(b.release(), [pool](KvBlock* raw) {
      std::unique_ptr<KvBlock> owned(raw);
      if (auto p = pool.lock()) {  // handle_return, reuse.rs:518-528
        ++p->available_;
        owned->return_tick = ++p->return_tick_;
        p->place(std::move(owned));
      }
    });
  }
  std::unordered_map<SequenceHash, std::unique_ptr<KvBlock>> lookup_;
  std::map<std::pair<uint32_t, uint64_t>, SequenceHash> priority_set_;  // (priority, return tick) -> sequence hash
  std::deque<std::unique_ptr<KvBlock>> uninitialized_;
  uint64_t return_tick_ = 0, total_ = 0, available_ = 0;
};

// In-flight blocks shared by sequence hash (reserved.rs:15-160): a block goes back to the pool when its last holder
// lets go.
using ReservedBlock = std::shared_ptr<KvBlock>;
class ReservedBlocks {
 public:
  explicit ReservedBlocks(size_t block_size) : block_size_(block_size) {}
  std::vector<ReservedBlock> match_sequence_hashes(const std::vector<SequenceHash>& hashes) {
    std::vector<ReservedBlock> out;
    for (SequenceHash h : hashes) {
      auto it = blocks_.find(h);
      ReservedBlock b = it == blocks_.end() ? nullptr : it->second.lock();
      if (!b) break;
      out.push_back(std::move(b));
    }
    return out;
  }
  std::vector<ReservedBlock> match_token_blocks(const std::vector<TokenBlock>& blocks) {
    std::vector<SequenceHash> h;
    for (const auto& b : blocks) h.push_back(b.sequence_hash);
    return match_sequence_hashes(h);
  }
  // reserved.rs:117-146: a second registration of the same hash drops the caller's duplicate (it returns to the pool)
  ReservedBlock register_block(PoolItem block) {
    if (block->token_block.tokens.size() != block_size_) raise("Block size mismatch");
    const SequenceHash h = block->token_block.sequence_hash;
    auto it = blocks_.find(h);
    if (it != blocks_.end())
      if (ReservedBlock live = it->second.lock()) return live;
    blocks_[h] = block;
    return block;
  }
  static long inflight_count(const ReservedBlock& b) { return b.use_count(); }

 private:
  size_t block_size_;
  std::unordered_map<SequenceHash, std::weak_ptr<KvBlock>> blocks_;
};

struct PrefillMatched {  // manager.rs:130-135
  std::vector<ReservedBlock> inflight_blocks;
  std::vector<TokenBlock> remaining_blocks;
  PartialTokenBlock tail_block;
};
struct PartialKvBlock {  // manager.rs:124-128
  PartialTokenBlock token_block;
  PoolItem kv_block;
};
struct PrefillOffload {  // manager.rs:137-142
  std::vector<ReservedBlock> inflight_blocks;
  std::vector<PoolItem> complete_prefill_blocks;
  PartialKvBlock tail_prefill_block;
  // cache slots of the whole request in sequence order: what the copy / transfer kernels take
  std::vector<int32_t> block_ids() const {
    std::vector<int32_t> ids;
    for (const auto& b : inflight_blocks) ids.push_back((int32_t)b->block_id);
    for (const auto& b : complete_prefill_blocks) ids.push_back((int32_t)b->block_id);
    ids.push_back((int32_t)tail_prefill_block.kv_block->block_id);
    return ids;
  }
};

class KvStorageManager {  // manager.rs:20-122
 public:
  explicit KvStorageManager(size_t block_size)
      : available_blocks(AvailableBlocks::create()), inflight_blocks(block_size), block_size_(block_size) {}
  PrefillMatched prepare_prefill_sequence(const std::vector<Token>& tokens) {
    TokenBlockSequence seq(tokens, block_size_, XXH3_SEED);  // crate::tokens: fixed seed 1337
    std::vector<TokenBlock> blocks = seq.blocks();
    std::vector<ReservedBlock> inflight = inflight_blocks.match_token_blocks(blocks);
    std::vector<SequenceHash> unmatched;
    for (size_t i = inflight.size(); i < blocks.size(); ++i) unmatched.push_back(blocks[i].sequence_hash);
    for (auto& b : available_blocks->match_blocks(unmatched)) inflight.push_back(inflight_blocks.register_block(b));
    std::vector<TokenBlock> remaining(blocks.begin() + inflight.size(), blocks.end());
    return PrefillMatched{std::move(inflight), std::move(remaining), seq.current_block()};
  }
  PrefillOffload prepare_prefill_offload(PrefillMatched matched) {
    const size_t need = matched.remaining_blocks.size() + 1;
    std::vector<PoolItem> reuse = available_blocks->take_blocks(need);
    if (reuse.size() != need) {
      const size_t got = reuse.size();
      reuse.clear();  // hands them back
      raise("expected " + std::to_string(need) + " blocks, got " + std::to_string(got));
    }
    std::vector<PoolItem> complete;
    for (auto& tb : matched.remaining_blocks) {
      PoolItem b = std::move(reuse.back());
      reuse.pop_back();
      b->update_token_block(std::move(tb));
      complete.push_back(std::move(b));
    }
    PartialKvBlock tail{std::move(matched.tail_block), std::move(reuse.back())};
    return PrefillOffload{std::move(matched.inflight_blocks), std::move(complete), std::move(tail)};
  }
  std::shared_ptr<AvailableBlocks> available_blocks;
  ReservedBlocks inflight_blocks;

 private:
  size_t block_size_;
};

}  // namespace kvdp
