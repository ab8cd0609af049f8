// host_types.hpp — C++ host mirror of the reference's integer layer (no CUDA in this header):
//   Sequence (src/engine/sequence.rs), Block / BlockManager (src/engine/block_manager.rs),
//   Scheduler (src/engine/scheduler.rs) and the runner's metadata builders (src/engine/model_runner.rs:159-300).
// Everything here is bit-exact against oracle/host.py (tests/test_host_parity.py): block ids, ref-counts, XXH64
// chained hashes, block tables, slot mappings, context lengths, cu_seqlens, scheduling and preemption order.
// Designed for the engine's step loop rather than translated: sequences live once in a slab owned by the
// scheduler (queues hold indices; the reference clones whole Sequence structs between queues), the free list is
// an intrusive doubly-linked list (O(1) removal; the reference's VecDeque::retain is O(num_blocks) per allocation),
// and the step metadata is written straight into reusable pinned-friendly flat arrays.
#pragma once

#include <cstdint>
#include <cstring>
#include <deque>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/nvllm_engine.h"

namespace nvllm {
namespace host {

// ---- XXH64 (third-party xxhash-rust 0.8 in the reference, call site block_manager.rs:122) ----------------
struct XXH64 {
  static constexpr uint64_t P1 = 0x9E3779B185EBCA87ull, P2 = 0xC2B2AE3D27D4EB4Full, P3 = 0x165667B19E3779F9ull,
                            P4 = 0x85EBCA77C2B2AE63ull, P5 = 0x27D4EB2F165667C5ull;
  static uint64_t rotl(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
  static uint64_t lane(uint64_t acc, uint64_t in) { return rotl(acc + in * P2, 31) * P1; }
  static uint64_t fold(uint64_t h, uint64_t v) { return (h ^ lane(0, v)) * P1 + P4; }
  // Hash of a stream of little-endian u64 words (all the path ever hashes: u64 prefix + i64 tokens).
  static uint64_t words(const uint64_t* w, size_t n, uint64_t seed) {
    size_t i = 0;
    uint64_t h;
    if (n >= 4) {
      uint64_t v1 = seed + P1 + P2, v2 = seed + P2, v3 = seed, v4 = seed - P1;
      for (; i + 4 <= n; i += 4) {
        v1 = lane(v1, w[i]); v2 = lane(v2, w[i + 1]); v3 = lane(v3, w[i + 2]); v4 = lane(v4, w[i + 3]);
      }
      h = rotl(v1, 1) + rotl(v2, 7) + rotl(v3, 12) + rotl(v4, 18);
      h = fold(h, v1); h = fold(h, v2); h = fold(h, v3); h = fold(h, v4);
    } else {
      h = seed + P5;
    }
    h += static_cast<uint64_t>(n) * 8;
    for (; i < n; ++i) h = rotl(h ^ lane(0, w[i]), 27) * P1 + P4;
    h ^= h >> 33; h *= P2; h ^= h >> 29; h *= P3; h ^= h >> 32;
    return h;
  }
};

// BlockManager::compute_hash (block_manager.rs:109-123)
inline uint64_t compute_hash(const int64_t* tokens, size_t n, bool has_prefix, uint64_t prefix) {
  std::vector<uint64_t> buf;
  buf.reserve(n + 1);
  if (has_prefix) buf.push_back(prefix);
  for (size_t i = 0; i < n; ++i) buf.push_back(static_cast<uint64_t>(tokens[i]));
  return XXH64::words(buf.data(), buf.size(), 0);
}

enum class Status : int32_t { Waiting = 0, Running = 1, Finished = 2, Preempted = 3 };

// ---- Sequence (sequence.rs:50-237) ----------------------------------------------------------------------
struct Sequence {
  int64_t seq_id = 0;
  Status status = Status::Waiting;
  std::vector<int64_t> token_ids;
  int64_t last_token = 0;
  int32_t num_prompt_tokens = 0;
  int32_t num_cached_tokens = 0;
  std::vector<int32_t> block_table;
  nvllm_sampling_params sp{};
  int32_t block_size = 256;  // follows Config (the reference hard-codes 256, sequence.rs:99)

  int32_t len() const { return static_cast<int32_t>(token_ids.size()); }
  int32_t num_completion_tokens() const { return len() - num_prompt_tokens; }
  int32_t num_blocks() const { return (len() + block_size - 1) / block_size; }           // :157-159
  int32_t last_block_num_tokens() const {                                                  // :167-174
    const int32_t r = len() % block_size;
    return (r == 0 && len() > 0) ? block_size : r;
  }
  void append_token(int64_t t) { token_ids.push_back(t); last_token = t; }                 // :150-154
  bool should_stop(int64_t eos) const {                                                    // :189-205
    if (num_completion_tokens() >= sp.max_tokens) return true;
    return !sp.ignore_eos && eos >= 0 && last_token == eos;
  }
};

// ---- BlockManager (block_manager.rs:69-331) -------------------------------------------------------------
class BlockManager {
 public:
  struct Block {
    int32_t ref_count = 0;
    bool has_hash = false;
    uint64_t hash = 0;
    std::vector<int64_t> token_ids;                              // kept only for hashed (shareable) blocks
    int32_t prev = -1, next = -1;                                // free-list links
    bool is_free_listed = false;
  };

  BlockManager(int32_t num_blocks, int32_t block_size) : block_size_(block_size), blocks_(num_blocks) {
    if (num_blocks <= 0 || block_size <= 0) throw std::invalid_argument("BlockManager: sizes must be positive");
    for (int32_t i = 0; i < num_blocks; ++i) push_back_free(i);  // free list = 0..n-1 (:96)
  }

  int32_t block_size() const { return block_size_; }
  int32_t num_blocks() const { return static_cast<int32_t>(blocks_.size()); }
  int32_t num_free() const { return free_count_; }
  int32_t num_used() const { return num_blocks() - free_count_; }
  size_t num_cached() const { return hash_to_block_.size(); }
  const Block& block(int32_t id) const { return blocks_[id]; }
  int32_t free_front() const { return free_head_; }
  int32_t free_next(int32_t id) const { return blocks_[id].next; }

  bool can_allocate(const Sequence& s) const { return free_count_ >= s.num_blocks(); }     // :152-154

  // :157-219 — chained-hash prefix reuse with token compare and the cache_miss latch.
  void allocate(Sequence& s) {
    if (!s.block_table.empty()) throw std::runtime_error("Sequence already has allocated blocks");
    if (!can_allocate(s)) throw std::runtime_error("Not enough free blocks to allocate sequence");
    bool has_prefix = false, cache_miss = false;
    uint64_t prefix = 0;
    const int32_t nb = s.num_blocks();
    for (int32_t bi = 0; bi < nb; ++bi) {
      const int32_t start = bi * block_size_;
      const int32_t n = std::min(block_size_, s.len() - start);
      const int64_t* toks = s.token_ids.data() + start;
      const bool full = n == block_size_;
      const uint64_t h = full ? compute_hash(toks, n, has_prefix, prefix) : 0;
      int32_t id;
      if (full) {
        auto it = hash_to_block_.find(h);
        if (it != hash_to_block_.end() && !cache_miss && same_tokens(blocks_[it->second], toks, n)) {
          id = it->second;
          s.num_cached_tokens += block_size_;                                              // :187
          if (!blocks_[id].is_free_listed) blocks_[id].ref_count += 1;                     // :189-192
          else take(id);                                                                   // :193-197
        } else {
          cache_miss = true;                                                               // :200,:205
          id = take_front();
          Block& b = blocks_[id];
          b.has_hash = true; b.hash = h; b.token_ids.assign(toks, toks + n);               // update (:229-231)
          hash_to_block_[h] = id;
        }
      } else {
        cache_miss = true;                                                                 // :210
        id = take_front();
      }
      s.block_table.push_back(id);                                                         // :214
      has_prefix = full;                                                                   // prefix_hash = current_hash (:215)
      prefix = h;
    }
  }

  void deallocate(Sequence& s) {                                                           // :240-252
    for (auto it = s.block_table.rbegin(); it != s.block_table.rend(); ++it) {
      Block& b = blocks_[*it];
      if (b.ref_count <= 0) throw std::runtime_error("Cannot remove reference from block with zero refs");
      if (--b.ref_count == 0) release(*it);
    }
    s.num_cached_tokens = 0;
    s.block_table.clear();
  }

  bool can_append(const Sequence& s) const {                                               // :255-262
    return (s.len() % block_size_ == 1) ? free_count_ > 0 : true;
  }

  void may_append(Sequence& s) {                                                           // :265-304
    if (s.block_table.empty()) throw std::runtime_error("Sequence has no allocated blocks");
    const int32_t last_id = s.block_table.back();
    Block& last = blocks_[last_id];
    const int32_t r = s.len() % block_size_;
    if (r == 1) {
      if (last.has_hash) {                                                                 // :276
        if (free_count_ == 0) throw std::runtime_error("No free blocks for append");
        s.block_table.push_back(take_front());                                             // :278-283
      }
    } else if (r == 0) {
      if (!last.has_hash) {                                                                // :287
        const int32_t start = (s.num_blocks() - 1) * block_size_;
        const size_t nt = s.block_table.size();
        const bool has_prefix = nt > 1 && blocks_[s.block_table[nt - 2]].has_hash;         // :289-294
        const uint64_t prefix = has_prefix ? blocks_[s.block_table[nt - 2]].hash : 0;
        const int64_t* toks = s.token_ids.data() + start;
        const uint64_t h = compute_hash(toks, block_size_, has_prefix, prefix);
        last.has_hash = true; last.hash = h; last.token_ids.assign(toks, toks + block_size_);
        hash_to_block_[h] = last_id;                                                       // :297-298
      }
    }
  }

 private:
  static bool same_tokens(const Block& b, const int64_t* toks, int32_t n) {
    return static_cast<int32_t>(b.token_ids.size()) == n &&
           std::memcmp(b.token_ids.data(), toks, sizeof(int64_t) * n) == 0;
  }
  void push_back_free(int32_t id) {
    Block& b = blocks_[id];
    b.prev = free_tail_; b.next = -1; b.is_free_listed = true;
    if (free_tail_ >= 0) blocks_[free_tail_].next = id; else free_head_ = id;
    free_tail_ = id;
    ++free_count_;
  }
  // allocate_block (:126-134): reset + unlink from the free list
  void take(int32_t id) {
    Block& b = blocks_[id];
    if (b.ref_count != 0) throw std::runtime_error("Block is not free");
    if (b.prev >= 0) blocks_[b.prev].next = b.next; else free_head_ = b.next;
    if (b.next >= 0) blocks_[b.next].prev = b.prev; else free_tail_ = b.prev;
    b.prev = b.next = -1; b.is_free_listed = false;
    --free_count_;
    b.ref_count = 1; b.has_hash = false; b.hash = 0; b.token_ids.clear();                  // Block::reset (:44-48)
  }
  int32_t take_front() {                                                                   // :223 front of the FIFO
    if (free_head_ < 0) throw std::runtime_error("No free blocks available");
    const int32_t id = free_head_;
    take(id);
    return id;
  }
  // deallocate_block (:137-149): back of the free list, drop the hash mapping if it still points here
  void release(int32_t id) {
    Block& b = blocks_[id];
    push_back_free(id);
    if (b.has_hash) {
      auto it = hash_to_block_.find(b.hash);
      if (it != hash_to_block_.end() && it->second == id) hash_to_block_.erase(it);
    }
  }

  int32_t block_size_;
  std::vector<Block> blocks_;
  std::unordered_map<uint64_t, int32_t> hash_to_block_;
  int32_t free_head_ = -1, free_tail_ = -1, free_count_ = 0;
};

}  // namespace host
}  // namespace nvllm
