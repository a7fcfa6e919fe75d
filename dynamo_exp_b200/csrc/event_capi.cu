// event_capi.cu — the engine-facing KV-event ABI, served from libkvdp.so.
//
// The reference ships these five symbols in libdynamo_llm_capi.so (lib/bindings/c/src/lib.rs:50-311); the patched vLLM
// loads whatever VLLM_KV_CAPI_PATH points at with ctypes and calls them from its block manager. What they do there:
// hash the tokens of every full block with seed 1337 (one compute_block_hash_for_seq call per block, :157-172), wrap
// the result in a KvCacheEvent (:175-248) and hand it to an unbounded channel whose consumer task publishes
// RouterEvent JSON on "<component subject>.kv_events" (kv_router/publisher.rs:40-78).
//
// Here the channel is a queue drained by one publisher thread. Whatever accumulated while the previous batch was in
// flight is hashed in ONE launch of the batched token-block hasher (kvdp_hash_token_blocks_host), so the per-event cost
// under load is a queue push; an idle queue is served immediately. Events leave in arrival order through the sinks
// chosen with kvdp_event_sink_*: a callback that receives (subject, RouterEvent JSON) — where a deployment hangs its
// transport (NATS in the reference; not rebuilt here), a prefix index that is updated in place, or, with neither, an
// in-memory outbox read with kvdp_event_poll.
#include <string.h>

#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {
using namespace kvdp;

constexpr uint64_t kSeed = 1337;         // kv_router/indexer.rs:64
constexpr size_t kOutboxMax = 1u << 16;  // events kept for kvdp_event_poll before the oldest are dropped

struct Pending {
  uint64_t event_id = 0;
  bool stored = false;
  bool has_parent = false;
  uint64_t parent = 0;
  std::vector<uint64_t> block_ids;  // stored: one per accepted block; removed: the hashes to drop
};

struct Publisher {
  std::mutex mu;
  std::condition_variable cv_work, cv_idle;
  std::thread th;
  bool initialised = false;  // dynamo_llm_init seen (and not shut down since)
  bool stop = false;
  std::string subject;
  int64_t worker_id = 0;
  uint32_t block_size = 0;
  int device = -1;
  // producer side
  std::vector<Pending> queue;
  std::vector<uint32_t> tokens;  // tokens of every accepted block of the queued stored events, back to back
  uint64_t submitted = 0, delivered = 0;
  // sinks
  kvdp_event_cb cb = nullptr;
  void* cb_user = nullptr;
  kvdp_index* ix = nullptr;
  int ix_worker = 0;
  std::deque<std::string> outbox;
  // counters (kvdp_event_stat)
  uint64_t n_batches = 0, n_blocks = 0, n_short_blocks = 0, n_parent_missing = 0, n_failed = 0, n_outbox_dropped = 0;
  uint32_t warned = 0;
};

Publisher& pub() {
  static Publisher* p = new Publisher();  // never destroyed: the thread may still be parked at process exit
  return *p;
}

void append_u64(std::string& s, uint64_t v) {
  char buf[24];
  int n = snprintf(buf, sizeof buf, "%llu", (unsigned long long)v);
  s.append(buf, (size_t)n);
}

// serde_json output of RouterEvent (indexer.rs:139-145, protocols.rs:79-121), compact, fields in declaration order
std::string to_json(int64_t worker_id, const Pending& e, const uint64_t* tokens_hash) {
  std::string s;
  s.reserve(96 + e.block_ids.size() * 64);
  char buf[32];
  int n = snprintf(buf, sizeof buf, "%lld", (long long)worker_id);
  s += "{\"worker_id\":";
  s.append(buf, (size_t)n);
  s += ",\"event\":{\"event_id\":";
  append_u64(s, e.event_id);
  if (e.stored) {
    s += ",\"data\":{\"stored\":{\"parent_hash\":";
    if (e.has_parent) append_u64(s, e.parent); else s += "null";
    s += ",\"blocks\":[";
    for (size_t i = 0; i < e.block_ids.size(); ++i) {
      if (i) s += ',';
      s += "{\"block_hash\":";
      append_u64(s, e.block_ids[i]);
      s += ",\"tokens_hash\":";
      append_u64(s, tokens_hash[i]);
      s += '}';
    }
    s += "]}}}}";
  } else {
    s += ",\"data\":{\"removed\":{\"block_hashes\":[";
    for (size_t i = 0; i < e.block_ids.size(); ++i) {
      if (i) s += ',';
      append_u64(s, e.block_ids[i]);
    }
    s += "]}}}}";
  }
  return s;
}

void publisher_loop() {
  Publisher& p = pub();
  std::vector<Pending> batch;
  std::vector<uint32_t> tokens;
  std::vector<int64_t> offsets, block_offsets;
  std::vector<uint64_t> block_hash, seq_hash;
  std::unique_lock<std::mutex> lk(p.mu);
  for (;;) {
    p.cv_work.wait(lk, [&] { return p.stop || !p.queue.empty(); });
    if (p.queue.empty()) break;  // stop requested and nothing left
    batch.clear();
    tokens.clear();
    batch.swap(p.queue);
    tokens.swap(p.tokens);
    const uint32_t bs = p.block_size;
    const int device = p.device;
    kvdp_event_cb cb = p.cb;
    void* cb_user = p.cb_user;
    kvdp_index* ix = p.ix;
    const int ix_worker = p.ix_worker;
    lk.unlock();

    // one request per block: block_hash[i] is the un-chained LocalBlockHash of block i
    const int64_t nb = bs ? (int64_t)(tokens.size() / bs) : 0;
    bool ok = true;
    if (nb > 0) {
      offsets.resize((size_t)nb + 1);
      for (int64_t i = 0; i <= nb; ++i) offsets[(size_t)i] = i * (int64_t)bs;
      block_offsets.resize((size_t)nb + 1);
      block_hash.resize((size_t)nb);
      seq_hash.resize((size_t)nb);
      int rc = kvdp_hash_token_blocks_host(tokens.data(), offsets.data(), nb, (int32_t)bs, kSeed, block_offsets.data(),
                                           block_hash.data(), seq_hash.data(), 0, device);
      if (rc != 0) {
        ok = false;
        fprintf(stderr, "kvdp: %zu KV events dropped, hashing their %lld blocks failed: %s\n", batch.size(),
                (long long)nb, kvdp_last_error());
      }
    }
    uint64_t parent_missing = 0;
    std::vector<std::string> for_outbox;
    if (ok) {
      size_t at = 0;
      for (const Pending& e : batch) {
        const uint64_t* th = e.stored && !e.block_ids.empty() ? &block_hash[at] : nullptr;
        if (e.stored) at += e.block_ids.size();
        if (ix) {
          int rc = e.stored ? kvdp_index_apply_stored(ix, ix_worker, e.has_parent, e.parent, e.block_ids.data(), th,
                                                      (int32_t)e.block_ids.size())
                            : kvdp_index_apply_removed(ix, ix_worker, e.block_ids.data(), (int32_t)e.block_ids.size());
          if (rc == 1) ++parent_missing;
          if (rc < 0) fprintf(stderr, "kvdp: prefix index rejected KV event %llu: %s\n",
                              (unsigned long long)e.event_id, kvdp_last_error());
        }
        if (cb || !ix) {
          std::string js = to_json(p.worker_id, e, th);
          if (cb) cb(cb_user, p.subject.c_str(), js.data(), js.size());
          else for_outbox.emplace_back(std::move(js));
        }
      }
    }
    lk.lock();
    p.n_batches += 1;
    p.n_blocks += (uint64_t)nb;
    p.n_parent_missing += parent_missing;
    if (!ok) p.n_failed += batch.size();
    for (auto& js : for_outbox) {
      if (p.outbox.size() >= kOutboxMax) {
        p.outbox.pop_front();
        ++p.n_outbox_dropped;
      }
      p.outbox.emplace_back(std::move(js));
    }
    p.delivered += batch.size();
    p.cv_idle.notify_all();
  }
}

int enqueue(Pending&& e, const uint32_t* toks, size_t ntok) {
  Publisher& p = pub();
  std::lock_guard<std::mutex> lk(p.mu);
  if (!p.initialised) return 1;
  if (ntok) p.tokens.insert(p.tokens.end(), toks, toks + ntok);
  p.queue.emplace_back(std::move(e));
  ++p.submitted;
  p.cv_work.notify_one();
  return 0;
}

}  // namespace

namespace kvdp {
// Called by kvdp_index_destroy: if the publisher's sink is this index, wait for the batch in flight and detach.
void event_sink_forget_index(const void* index) {
  Publisher& p = pub();
  std::unique_lock<std::mutex> lk(p.mu);
  if (p.ix != index) return;
  p.cv_idle.wait(lk, [&] { return p.delivered == p.submitted; });
  p.ix = nullptr;
}
}  // namespace kvdp

extern "C" {

// lib.rs:50-113. The reference keeps its publisher in a process-wide OnceCell, so a second init returns OK and changes
// nothing; same here (until dynamo_llm_shutdown, after which a new init starts a new publisher).
dynamo_llm_result_t dynamo_llm_init(const char* namespace_c_str, const char* component_c_str, int64_t worker_id,
                                    uint32_t kv_block_size) {
  if (!namespace_c_str || !component_c_str) {
    fprintf(stderr, "kvdp: dynamo_llm_init: NULL namespace or component\n");
    return DYNAMO_LLM_ERR;
  }
  if (kv_block_size == 0) {
    fprintf(stderr, "kvdp: dynamo_llm_init: kv_block_size must be positive\n");
    return DYNAMO_LLM_ERR;
  }
  Publisher& p = pub();
  std::lock_guard<std::mutex> lk(p.mu);
  if (p.initialised) return DYNAMO_LLM_OK;
  p.subject = std::string("namespace.") + namespace_c_str + ".component." + component_c_str + ".kv_events";
  p.worker_id = worker_id;
  p.block_size = kv_block_size;
  if (cudaGetDevice(&p.device) != cudaSuccess) {
    p.device = -1;
    cudaGetLastError();
  }
  p.stop = false;
  p.initialised = true;
  p.th = std::thread(publisher_loop);
  return DYNAMO_LLM_OK;
}

// lib.rs:115-128: ERR when nothing was initialised. Queued events are delivered before this returns.
dynamo_llm_result_t dynamo_llm_shutdown(void) {
  Publisher& p = pub();
  std::thread th;
  {
    std::lock_guard<std::mutex> lk(p.mu);
    if (!p.initialised) {
      fprintf(stderr, "Runtime not initialized\n");
      return DYNAMO_LLM_ERR;
    }
    p.initialised = false;
    p.stop = true;
    p.cv_work.notify_all();
    th.swap(p.th);
  }
  if (th.joinable()) th.join();
  return DYNAMO_LLM_OK;
}

dynamo_llm_result_t dynamo_llm_load_publisher_create(void) { return DYNAMO_LLM_OK; }  // lib.rs:130-133

// lib.rs:175-228,260-295. Blocks are taken in order until one whose token count is not kv_block_size; that block and
// everything after it are left out, and the event (possibly with no blocks) is still published.
dynamo_llm_result_t dynamo_kv_event_publish_stored(uint64_t event_id, const uint32_t* token_ids,
                                                   const size_t* num_block_tokens, const uint64_t* block_ids,
                                                   size_t num_blocks, const uint64_t* parent_hash, uint64_t lora_id) {
  (void)lora_id;  // unused by the reference too (:162)
  Publisher& p = pub();
  uint32_t bs;
  {
    std::lock_guard<std::mutex> lk(p.mu);
    if (!p.initialised) {  // the reference unwraps an empty OnceCell here and aborts the process
      fprintf(stderr, "kvdp: dynamo_kv_event_publish_stored called before dynamo_llm_init\n");
      return DYNAMO_LLM_ERR;
    }
    bs = p.block_size;
  }
  if (num_blocks && (!token_ids || !num_block_tokens || !block_ids)) {
    fprintf(stderr, "kvdp: dynamo_kv_event_publish_stored: NULL array\n");
    return DYNAMO_LLM_ERR;
  }
  Pending e;
  e.event_id = event_id;
  e.stored = true;
  e.has_parent = parent_hash != nullptr;
  e.parent = parent_hash ? *parent_hash : 0;
  size_t ntok = 0;
  for (size_t b = 0; b < num_blocks; ++b) {
    if (num_block_tokens[b] != bs) {
      std::lock_guard<std::mutex> lk(p.mu);
      ++p.n_short_blocks;
      if (p.warned < 3) {  // :191-205 warns three times, then stays quiet
        ++p.warned;
        fprintf(stderr, "Block not published. Block size must be %u tokens to be published. Block size is: %zu\n", bs,
                num_block_tokens[b]);
      }
      break;
    }
    ntok += bs;
    e.block_ids.push_back(block_ids[b]);
  }
  return enqueue(std::move(e), token_ids, ntok) ? DYNAMO_LLM_ERR : DYNAMO_LLM_OK;
}

// lib.rs:230-248,297-311
dynamo_llm_result_t dynamo_kv_event_publish_removed(uint64_t event_id, const uint64_t* block_ids, size_t num_blocks) {
  if (num_blocks && !block_ids) {
    fprintf(stderr, "kvdp: dynamo_kv_event_publish_removed: NULL array\n");
    return DYNAMO_LLM_ERR;
  }
  Pending e;
  e.event_id = event_id;
  e.stored = false;
  e.block_ids.assign(block_ids, block_ids + num_blocks);
  if (enqueue(std::move(e), nullptr, 0)) {
    fprintf(stderr, "kvdp: dynamo_kv_event_publish_removed called before dynamo_llm_init\n");
    return DYNAMO_LLM_ERR;
  }
  return DYNAMO_LLM_OK;
}

// ---------------------------------------------------------------- sinks and draining (no reference counterpart)
int kvdp_event_sink_callback(kvdp_event_cb cb, void* user) {
  Publisher& p = pub();
  std::lock_guard<std::mutex> lk(p.mu);
  p.cb = cb;
  p.cb_user = user;
  return 0;
}

int kvdp_event_sink_index(kvdp_index* ix, int32_t worker_slot) {
  if (worker_slot < 0 || worker_slot >= 64) return fail(KVDP_EINVAL, "kvdp_event_sink_index: worker slot out of range");
  Publisher& p = pub();
  std::unique_lock<std::mutex> lk(p.mu);
  p.cv_idle.wait(lk, [&] { return p.delivered == p.submitted; });  // nothing in flight still points at the old index
  p.ix = ix;
  p.ix_worker = worker_slot;
  return 0;
}

int kvdp_event_flush(void) {
  Publisher& p = pub();
  std::unique_lock<std::mutex> lk(p.mu);
  const uint64_t target = p.submitted;
  p.cv_idle.wait(lk, [&] { return p.delivered >= target; });
  return 0;
}

int kvdp_event_poll(char* buf, size_t cap, size_t* len) {
  if (!len) return fail(KVDP_EINVAL, "kvdp_event_poll: NULL len");
  Publisher& p = pub();
  std::lock_guard<std::mutex> lk(p.mu);
  if (p.outbox.empty()) {
    *len = 0;
    return 0;
  }
  const std::string& js = p.outbox.front();
  *len = js.size();
  if (!buf || cap < js.size()) return fail(KVDP_ERANGE, "kvdp_event_poll: next event needs %zu bytes", js.size());
  memcpy(buf, js.data(), js.size());
  p.outbox.pop_front();
  return 1;
}

int64_t kvdp_event_stat(int32_t what) {
  Publisher& p = pub();
  std::lock_guard<std::mutex> lk(p.mu);
  switch (what) {
    case 0: return (int64_t)p.submitted;
    case 1: return (int64_t)p.delivered;
    case 2: return (int64_t)p.n_batches;
    case 3: return (int64_t)p.n_blocks;
    case 4: return (int64_t)p.n_short_blocks;
    case 5: return (int64_t)p.n_parent_missing;
    case 6: return (int64_t)p.n_failed;
    case 7: return (int64_t)p.n_outbox_dropped;
    case 8: return (int64_t)p.outbox.size();
    default: return -1;
  }
}

}  // extern "C"
