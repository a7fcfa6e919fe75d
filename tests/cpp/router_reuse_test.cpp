// C++ host-layer test for include/kvdp_router.hpp.
//   pool part (no GPU):  the reference's pool tests, kv/reuse.rs:684-1062 and kv/reserved.rs:161-242, with hand-made
//                        sequence hashes (the pool never looks inside a hash)
//   GPU part:            RadixTree scenario of kv_router/indexer.rs:949-1150, sharded indexer, the event ABI feeding an
//                        index (test_kv_bindings.py:66-99), KvStorageManager on GPU-hashed tokens
// `router_reuse_test pool` runs the first part only. Prints "CPP ROUTER/REUSE OK" and exits 0 when every check holds.
#include <cstdio>
#include <cstring>

#include "kvdp_router.hpp"

using namespace kvdp;

static int failures = 0;
#define CHECK(cond)                                                           \
  do {                                                                        \
    if (!(cond)) {                                                            \
      std::fprintf(stderr, "CHECK failed at line %d: %s\n", __LINE__, #cond); \
      ++failures;                                                             \
    }                                                                         \
  } while (0)

// create_blocks (reuse.rs:751-766) with synthetic chained hashes: block i of a sequence gets base + i
static std::vector<KvBlock> make_blocks(const std::vector<Token>& values, size_t bs, uint64_t base, uint32_t prio = 0) {
  std::vector<KvBlock> out;
  for (size_t i = 0; i < values.size() / bs; ++i) {
    KvBlock b;
    b.token_block.tokens.assign(values.begin() + i * bs, values.begin() + (i + 1) * bs);
    b.token_block.salt_hash = 1337;
    b.token_block.block_hash = 7000 + base + i;
    b.token_block.sequence_hash = base + i;
    if (i) b.token_block.parent_sequence_hash = base + i - 1;
    b.priority = prio;
    out.push_back(b);
  }
  return out;
}

static std::vector<Token> first_tokens(const std::vector<PoolItem>& items) {
  std::vector<Token> t;
  for (const auto& i : items) t.push_back(i->token_block.tokens[0]);
  return t;
}

static std::shared_ptr<AvailableBlocks> two_sequences(uint32_t p1, uint32_t p2) {
  auto pool = AvailableBlocks::create();
  auto b1 = make_blocks({1, 2, 3, 4}, 2, 100, p1), b2 = make_blocks({5, 6, 7, 8}, 2, 200, p2);
  for (auto it = b2.rbegin(); it != b2.rend(); ++it) pool->insert(*it);
  for (auto it = b1.rbegin(); it != b1.rend(); ++it) pool->insert(*it);
  return pool;
}

static void test_pool() {
  {  // reuse.rs:768-808
    auto pool = AvailableBlocks::create();
    for (auto& b : make_blocks({1, 2, 3, 4}, 2, 100)) pool->insert(b);
    pool->fence();
    CHECK(pool->total_blocks() == 2 && pool->available_blocks() == 2);
    auto matched = pool->match_blocks({100, 101});
    CHECK(matched.size() == 2 && pool->total_blocks() == 2 && pool->available_blocks() == 0);
    CHECK(matched[0]->token_block.sequence_hash == 100 && matched[1]->token_block.sequence_hash == 101);
    matched.pop_back();
    matched.pop_back();  // dropped: back in the pool
    CHECK(pool->total_blocks() == 2 && pool->available_blocks() == 2);
    auto partial = pool->match_blocks({100, 999, 101});
    CHECK(partial.size() == 1 && partial[0]->token_block.sequence_hash == 100);
  }
  CHECK((first_tokens(two_sequences(1, 1)->take_blocks(4)) == std::vector<Token>{7, 5, 3, 1}));  // reuse.rs:810-845
  CHECK((first_tokens(two_sequences(1, 2)->take_blocks(4)) == std::vector<Token>{3, 1, 7, 5}));  // reuse.rs:847-882
  {  // reuse.rs:884-935
    auto pool = two_sequences(1, 1);
    pool->update_multiple({UpdateBlock{200, 2}, UpdateBlock{201, 2}});
    CHECK((first_tokens(pool->take_blocks(4)) == std::vector<Token>{3, 1, 7, 5}));
  }
  {  // reuse.rs:937-1030
    auto pool = two_sequences(1, 1);
    pool->reset_all();
    CHECK(pool->match_blocks({200, 201}).empty() && pool->available_blocks() == 4);
    CHECK(pool->take_blocks(10).size() == 4);
    pool = two_sequences(1, 1);
    pool->reset({200, 201});
    CHECK(pool->match_blocks({200, 201}).empty());
    CHECK(pool->match_blocks({100, 101}).size() == 2);
  }
  {  // duplicates and hash 0 wait in the un-initialised queue and are handed out first (reuse.rs:317-323, 400-420)
    auto pool = AvailableBlocks::create();
    KvBlock a = make_blocks({1, 2}, 2, 50)[0], b = make_blocks({1, 2}, 2, 50)[0];
    a.block_id = 1;
    b.block_id = 2;
    pool->insert(a);
    pool->insert(b);
    pool->insert(KvBlock{});
    auto first = pool->take_blocks(1);
    CHECK(first.size() == 1 && first[0]->block_id == 2);
  }
  {  // reserved.rs:66-150, tests :161-242
    auto pool = AvailableBlocks::create();
    for (auto& b : make_blocks({1, 2, 3, 4}, 2, 300)) pool->insert(b);
    ReservedBlocks reserved(2);
    auto items = pool->match_blocks({300, 301});
    ReservedBlock r0 = reserved.register_block(items[0]), r1 = reserved.register_block(items[1]);
    items.clear();
    CHECK(pool->available_blocks() == 0 && ReservedBlocks::inflight_count(r0) == 1);
    auto again = reserved.match_sequence_hashes({300, 301, 302});
    CHECK(again.size() == 2 && ReservedBlocks::inflight_count(r0) == 2);  // shared, not copied
    pool->insert(make_blocks({1, 2}, 2, 300)[0]);
    ReservedBlock r0b;
    {
      auto dup = pool->take_blocks(1);
      r0b = reserved.register_block(dup[0]);  // gets the live block; the duplicate returns to the pool at scope end
    }
    CHECK(r0b == r0 && ReservedBlocks::inflight_count(r0) == 3 && pool->available_blocks() == 1);
    r0.reset();
    r0b.reset();
    again[0].reset();
    CHECK(reserved.match_sequence_hashes({300}).empty());  // last holder gone: back in the pool
    CHECK(pool->available_blocks() == 2);
    bool threw = false;
    try {
      auto pool2 = AvailableBlocks::create();
      pool2->insert(make_blocks({9}, 1, 900)[0]);
      reserved.register_block(pool2->take_blocks(1)[0]);
    } catch (const Error& e) {
      threw = std::strstr(e.what(), "Block size mismatch") != nullptr;
    }
    CHECK(threw);
    r1.reset();
    again[1].reset();
    CHECK(pool->available_blocks() == 3);
  }
}

// make_blocks of indexer.rs:900-908: tokens_hash = i, block_hash = i * 100
static RouterEvent store_event(WorkerId w, uint64_t id, std::vector<uint64_t> hashes,
                               std::optional<uint64_t> parent = std::nullopt) {
  RouterEvent e;
  e.worker_id = w;
  e.event.event_id = id;
  e.event.stored = KvCacheStoreData{parent, {}};
  for (uint64_t h : hashes) e.event.stored->blocks.push_back({h * 100, h});
  return e;
}
static RouterEvent remove_event(WorkerId w, uint64_t id, std::vector<uint64_t> hashes) {
  RouterEvent e;
  e.worker_id = w;
  e.event.event_id = id;
  e.event.removed = KvCacheRemoveData{};
  for (uint64_t h : hashes) e.event.removed->block_hashes.push_back(h * 100);
  return e;
}

static void test_index() {  // indexer.rs:949-1150
  RadixTree trie(1024);
  const WorkerId w1 = 0, w2 = 1;
  trie.apply_event(store_event(w1, 1, {1, 2, 3}));
  CHECK(trie.find_matches({1, 2, 3}).scores[w1] == 3);
  trie.apply_event(store_event(w2, 1, {1, 4, 5}));
  auto s = trie.find_matches({1, 2, 3}).scores;
  CHECK(s[w1] == 3 && s[w2] == 1);
  trie.apply_event(remove_event(w2, 2, {5}));
  trie.apply_event(remove_event(w2, 3, {4}));
  trie.apply_event(store_event(w2, 4, {2, 6, 7}, 100));
  s = trie.find_matches({1, 2, 3}).scores;
  CHECK(s[w1] == 3 && s[w2] == 2);
  CHECK(trie.stat(2, (uint64_t)trie.slot_of(w1, false)) == 3 && trie.stat(2, (uint64_t)trie.slot_of(w2, false)) == 4);
  trie.remove_worker(w1);  // indexer.rs:1152-1175
  s = trie.find_matches({1, 2, 3}).scores;
  CHECK(s.count(w1) == 0 && s[w2] == 2);
  auto batch = trie.find_matches_batch({{1, 2, 3}, {9}, {1, 2, 6, 7}});
  CHECK(batch[0].scores[w2] == 2 && batch[1].scores.empty() && batch[2].scores[w2] == 4);
  // early exit (indexer.rs:1177-1200): stop at the first block only one worker holds
  RadixTree t2(256);
  t2.apply_event(store_event(0, 0, {1, 2, 3}));
  t2.apply_event(store_event(1, 0, {1}));
  CHECK(t2.find_matches({1, 2, 3}, true).scores[0] == 2 && t2.find_matches({1, 2, 3}, true).scores[1] == 1);

  {  // indexer.rs:1305-1363 (test_frequency), clock passed in
    RadixTree f(256, -1, 0.050);
    f.apply_event(store_event(0, 0, {1, 2, 3, 4}));
    CHECK(f.find_matches({1, 2, 3, 4}, false, 0.005).frequencies.empty());
    CHECK((f.find_matches({1, 2, 3, 4}, false, 0.0051).frequencies == std::vector<size_t>{1, 1, 1, 1}));
    CHECK(f.find_matches({1, 2, 3, 4}, false, 0.105).frequencies.empty());
    CHECK((f.find_matches({1, 2, 3, 4}, false, 0.1051).frequencies == std::vector<size_t>{1, 1, 1, 1}));
    CHECK((f.find_matches({1, 2, 3}, false, 0.1052).frequencies == std::vector<size_t>{2, 2, 2}));
    CHECK((f.find_matches({1, 2, 3, 4}, false, 0.1053).frequencies == std::vector<size_t>{3, 3, 3, 2}));
  }

  KvIndexerSharded sharded(3, 4, 1024);
  for (WorkerId w = 0; w < 100; ++w) sharded.apply_event(store_event(w, 0, {1, 2, (uint64_t)(10 + w)}));
  CHECK((sharded.worker_counts() == std::vector<size_t>{34, 33, 33}));
  auto all = sharded.find_matches({1, 2, 55});
  CHECK(all.scores.size() == 100 && all.scores[45] == 3 && all.scores[7] == 2);
  sharded.remove_worker(45);
  CHECK(sharded.find_matches({1, 2, 55}).scores.count(45) == 0);
}

static void test_publisher_and_manager() {
  // test_kv_bindings.py:66-99 through the C ABI: store one block, the worker scores 1; remove it, no scores
  const size_t bs = 32;
  KvIndexer indexer(bs, 1024);
  KvEventPublisher pub("kv_test", "event", 233, bs);
  pub.attach_index(&indexer.tree());
  std::vector<Token> tok(bs, 3);
  CHECK(indexer.find_matches_for_request(tok).scores.empty());
  pub.publish_stored(0, tok, {bs}, {0}, 0);
  pub.flush();
  auto sc = indexer.find_matches_for_request(tok).scores;
  CHECK(sc.size() == 1 && sc[233] == 1);
  pub.publish_removed(1, {0});
  pub.flush();
  CHECK(indexer.find_matches_for_request(tok).scores.empty());
  pub.attach_index(nullptr);
  pub.publish_stored(2, {1, 2, 3, 4}, {4}, {77}, 0, 5);  // short block for bs = 32: published with no blocks
  pub.flush();
  auto js = pub.poll();
  CHECK(js && *js == "{\"worker_id\":233,\"event\":{\"event_id\":2,\"data\":{\"stored\":{\"parent_hash\":5,\"blocks\":[]}}}}");
  CHECK(!pub.poll());
  pub.shutdown();

  // manager.rs:37-122 on GPU-hashed tokens
  const size_t b4 = 4;
  KvStorageManager mgr(b4);
  for (int i = 0; i < 32; ++i) {
    KvBlock b;
    b.block_id = i;
    mgr.available_blocks->insert(b);
  }
  std::vector<Token> toks;
  for (Token t = 100; t < 100 + 4 * b4 + 2; ++t) toks.push_back(t);
  auto matched = mgr.prepare_prefill_sequence(toks);
  CHECK(matched.inflight_blocks.empty() && matched.remaining_blocks.size() == 4 && matched.tail_block.tokens().size() == 2);
  auto off = mgr.prepare_prefill_offload(std::move(matched));
  auto ids = off.block_ids();
  CHECK(ids.size() == 5 && std::set<int32_t>(ids.begin(), ids.end()).size() == 5);
  CHECK(mgr.available_blocks->available_blocks() == 27);
  std::vector<ReservedBlock> registered;
  for (auto& b : off.complete_prefill_blocks) registered.push_back(mgr.inflight_blocks.register_block(b));
  off.complete_prefill_blocks.clear();
  std::vector<Token> toks2(toks.begin(), toks.begin() + 4 * b4);
  for (Token t : {7u, 7u, 7u, 7u, 9u}) toks2.push_back(t);
  auto m2 = mgr.prepare_prefill_sequence(toks2);
  CHECK(m2.inflight_blocks.size() == 4 && m2.remaining_blocks.size() == 1 && m2.tail_block.tokens().size() == 1);
  for (size_t i = 0; i < 4 && i < m2.inflight_blocks.size(); ++i) CHECK(m2.inflight_blocks[i]->block_id == ids[i]);
  registered.clear();
  m2.inflight_blocks.clear();
  CHECK(mgr.available_blocks->available_blocks() == 31);  // only the first request's tail block is still out
  auto m3 = mgr.prepare_prefill_sequence(toks);
  CHECK(m3.inflight_blocks.size() == 4 && m3.remaining_blocks.empty());
  bool threw = false;
  try {
    KvStorageManager tiny(b4);
    tiny.prepare_prefill_offload(tiny.prepare_prefill_sequence(toks));
  } catch (const Error& e) {
    threw = std::strstr(e.what(), "expected 5 blocks, got 0") != nullptr;
  }
  CHECK(threw);
}

int main(int argc, char** argv) {
  const bool pool_only = argc > 1 && !std::strcmp(argv[1], "pool");
  try {
    test_pool();
    if (!pool_only) {
      int n = 0;
      if (kvdp_device_count(&n) != 0 || n < 1) {
        std::fprintf(stderr, "no CUDA device: %s\n", kvdp_last_error());
        return 2;
      }
      test_index();
      test_publisher_and_manager();
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "unexpected exception: %s\n", e.what());
    return 1;
  }
  if (failures) {
    std::fprintf(stderr, "%d check(s) failed\n", failures);
    return 1;
  }
  std::printf(pool_only ? "CPP POOL OK\n" : "CPP ROUTER/REUSE OK\n");
  return 0;
}
