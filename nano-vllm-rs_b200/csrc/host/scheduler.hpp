// scheduler.hpp — Scheduler (reference src/engine/scheduler.rs:14-334) + the step-metadata builders of
// ModelRunner (src/engine/model_runner.rs:159-300).  Pure host code; see host_types.hpp for the design notes.
#pragma once

#include <algorithm>

#include "host_types.hpp"

namespace nvllm {
namespace host {

// Flat, reusable arrays behind an nvllm_batch (the explicit replacement of the reference's global Context).
struct BatchArrays {
  std::vector<int64_t> seq_ids, input_ids, positions;
  std::vector<int32_t> slot_mapping, cu_q, cu_k, context_lens, block_tables;
  std::vector<float> temperatures, top_p;
  std::vector<int32_t> top_k;
  std::vector<int32_t> seq_index;  // slab indices of the scheduled sequences (postprocess order)
};

class Scheduler {
 public:
  explicit Scheduler(const nvllm_config& c)
      : max_num_seqs_(c.max_num_seqs), max_num_batched_tokens_(c.max_num_batched_tokens),
        max_model_len_(c.max_model_len), eos_(c.eos_token_id), block_size_(c.kvcache_block_size),
        bm_(c.num_kvcache_blocks > 0 ? c.num_kvcache_blocks : 1000, c.kvcache_block_size) {}  // :71-74

  BlockManager& block_manager() { return bm_; }
  const BlockManager& block_manager() const { return bm_; }
  bool is_finished() const { return waiting_.empty() && running_.empty(); }                   // :88-90
  // Sequence by id, or nullptr once it has been released (or never existed).
  const Sequence* find(int64_t seq_id) const {
    auto it = id_to_slot_.find(seq_id);
    return it == id_to_slot_.end() ? nullptr : &seqs_[it->second];
  }

  // Admission limits the reference never checks (it carries max_model_len, config.rs:27,60, but nothing reads it): the
  // runner sizes its block tables and RoPE table for max_model_len tokens, so a request that could outgrow them is
  // bounded HERE, before any scheduler state changes.  A prompt longer than max_model_len (or than one prefill batch,
  // which would never be admitted: scheduler.rs:136) is refused; max_tokens is clamped so the sequence stops with
  // len == max_model_len + 1 at the latest (the last sampled token is never fed back, so it needs no KV slot).
  int64_t add_sequence(const int64_t* prompt, int32_t n, const nvllm_sampling_params& sp) {   // :93-98
    if (max_model_len_ > 0 && n > max_model_len_)
      throw std::invalid_argument("prompt of " + std::to_string(n) + " tokens exceeds max_model_len " + std::to_string(max_model_len_));
    if (n > max_num_batched_tokens_)
      throw std::invalid_argument("prompt of " + std::to_string(n) + " tokens exceeds max_num_batched_tokens " +
                                  std::to_string(max_num_batched_tokens_));
    Sequence s;
    s.seq_id = next_seq_id_++;  // per-scheduler counter (reference: process-global AtomicU64)
    s.token_ids.assign(prompt, prompt + n);
    s.last_token = n > 0 ? prompt[n - 1] : 0;
    s.num_prompt_tokens = n;
    s.sp = sp;
    if (max_model_len_ > 0 && s.sp.max_tokens > max_model_len_ - n + 1) s.sp.max_tokens = max_model_len_ - n + 1;
    s.block_size = block_size_;
    int32_t slot;
    if (!free_slots_.empty()) { slot = free_slots_.back(); free_slots_.pop_back(); seqs_[slot] = std::move(s); }
    else { slot = static_cast<int32_t>(seqs_.size()); seqs_.push_back(std::move(s)); }
    id_to_slot_[seqs_[slot].seq_id] = slot;
    waiting_.push_back(slot);
    ++total_sequences_;
    return seqs_[slot].seq_id;
  }

  // Drops a FINISHED sequence's storage and recycles its slab slot (the reference hands finished sequences to the caller
  // and forgets them, llm_engine.rs:186-193; without this a long-running engine grows without bound).
  bool release(int64_t seq_id) {
    auto it = id_to_slot_.find(seq_id);
    if (it == id_to_slot_.end() || seqs_[it->second].status != Status::Finished) return false;
    Sequence& s = seqs_[it->second];
    std::vector<int64_t>().swap(s.token_ids);
    std::vector<int32_t>().swap(s.block_table);
    free_slots_.push_back(it->second);
    id_to_slot_.erase(it);
    return true;
  }

  // Scheduler::schedule (:103-116): prefill first, else decode.  Fills `out` (pointers into arrays_).
  void schedule(nvllm_batch* out) {
    arrays_.seq_index.clear();
    if (try_schedule_prefill()) {
      ++prefill_batches_;
      build_prefill(out);
      return;
    }
    try_schedule_decode();
    ++decode_batches_;
    build_decode(out);
  }

  // Scheduler::postprocess (:234-257)
  void postprocess(const int64_t* tokens, int32_t n) {
    if (n != static_cast<int32_t>(arrays_.seq_index.size()))
      throw std::runtime_error("Mismatch between sequences and token_ids length");
    for (int32_t i = 0; i < n; ++i) {
      const int32_t si = arrays_.seq_index[i];
      Sequence& s = seqs_[si];
      s.append_token(tokens[i]);
      if (s.should_stop(eos_)) {
        s.status = Status::Finished;
        bm_.deallocate(s);
        running_.erase(std::remove(running_.begin(), running_.end(), si), running_.end());    // :260-262
        ++finished_;
      }
    }
    arrays_.seq_index.clear();
  }

  void stats(nvllm_sched_stats* o) const {
    o->total_sequences = total_sequences_; o->waiting = static_cast<int64_t>(waiting_.size());
    o->running = static_cast<int64_t>(running_.size()); o->finished = finished_; o->preemptions = preemptions_;
    o->prefill_batches = prefill_batches_; o->decode_batches = decode_batches_;
    o->total_blocks = bm_.num_blocks(); o->free_blocks = bm_.num_free(); o->used_blocks = bm_.num_used();
    o->cached_blocks = static_cast<int64_t>(bm_.num_cached());
  }

 private:
  // :119-168
  bool try_schedule_prefill() {
    int32_t num_seqs = 0, num_batched = 0;
    std::vector<int32_t>& sched = arrays_.seq_index;
    while (!waiting_.empty()) {
      Sequence& s = seqs_[waiting_.front()];
      if (num_seqs >= max_num_seqs_) break;                                                   // :131
      const int32_t seq_tokens = s.len() - s.num_cached_tokens;                               // :135 (before allocate)
      if (num_batched + seq_tokens > max_num_batched_tokens_) break;                          // :136 strict >
      if (!bm_.can_allocate(s)) break;                                                        // :141
      sched.push_back(waiting_.front());
      waiting_.pop_front();
      bm_.allocate(s);                                                                        // :149
      ++num_seqs;
      num_batched += seq_tokens;
      s.status = Status::Running;
    }
    for (int32_t si : sched) running_.push_back(si);                                          // :163-165
    return !sched.empty();
  }

  // :171-223
  void try_schedule_decode() {
    std::vector<int32_t>& sched = arrays_.seq_index;
    std::vector<int32_t> resched;
    int32_t num_seqs = 0;
    while (!running_.empty()) {
      const int32_t si = running_.front();
      running_.pop_front();                                                                   // :178
      if (num_seqs >= max_num_seqs_) { resched.push_back(si); continue; }                     // :179-182
      Sequence& s = seqs_[si];
      bool preempted_self = false;
      while (!bm_.can_append(s)) {                                                            // :185
        if (!running_.empty()) { preempt(running_.back()); running_.pop_back(); }             // :186-188 tail victim
        else if (!sched.empty()) { preempt(sched.back()); sched.pop_back(); }  // :189-192
        else { preempt(si); preempted_self = true; break; }                                   // :193-196
      }
      if (!preempted_self && bm_.can_append(s)) {                                             // :201 (intent)
        ++num_seqs;
        bm_.may_append(s);                                                                    // :203
        sched.push_back(si);
      }
    }
    for (auto it = resched.rbegin(); it != resched.rend(); ++it) running_.push_front(*it);    // :209-211
    for (auto it = sched.rbegin(); it != sched.rend(); ++it) running_.push_front(*it);        // :214-216
    if (sched.empty()) throw std::runtime_error("No sequences could be scheduled for decode"); // :218-220
  }

  void preempt(int32_t si) {                                                                  // :226-231
    Sequence& s = seqs_[si];
    s.status = Status::Preempted;
    bm_.deallocate(s);
    waiting_.push_front(si);
    ++preemptions_;
  }

  void fill_common(nvllm_batch* out, bool is_prefill) {
    BatchArrays& a = arrays_;
    const int32_t n = static_cast<int32_t>(a.seq_index.size());
    a.seq_ids.resize(n);
    a.temperatures.resize(n); a.top_k.resize(n); a.top_p.resize(n);
    for (int32_t i = 0; i < n; ++i) {                                                         // model_runner.rs:137-146
      const Sequence& s = seqs_[a.seq_index[i]];
      a.seq_ids[i] = s.seq_id;
      a.temperatures[i] = s.sp.temperature;
      a.top_k[i] = s.sp.top_k;
      a.top_p[i] = s.sp.top_p;
    }
    std::memset(out, 0, sizeof(*out));
    out->is_prefill = is_prefill ? 1 : 0;
    out->num_seqs = n;
    out->seq_ids = a.seq_ids.data();
    out->temperatures = a.temperatures.data();
    out->top_k = a.top_k.data();
    out->top_p = a.top_p.data();
  }

  void fill_tables(nvllm_batch* out) {                                                        // model_runner.rs:280-293
    BatchArrays& a = arrays_;
    size_t width = 1;
    for (int32_t si : a.seq_index) width = std::max(width, seqs_[si].block_table.size());
    a.block_tables.assign(a.seq_index.size() * width, -1);
    for (size_t i = 0; i < a.seq_index.size(); ++i) {
      const auto& t = seqs_[a.seq_index[i]].block_table;
      std::copy(t.begin(), t.end(), a.block_tables.begin() + i * width);
    }
    out->max_blocks = static_cast<int32_t>(width);
    out->block_tables = a.block_tables.data();
    out->has_block_tables = 1;
  }

  // prepare_prefill_inputs + create_prefill_context (model_runner.rs:172-193,222-263) with the intended slot
  // mapping and cached-prefix handling (oracle/host.py prepare_prefill).
  void build_prefill(nvllm_batch* out) {
    fill_common(out, true);
    BatchArrays& a = arrays_;
    a.input_ids.clear(); a.positions.clear(); a.slot_mapping.clear();
    a.cu_q.assign(1, 0); a.cu_k.assign(1, 0);
    int32_t max_q = 0, max_k = 0;
    bool any_cached = false;
    for (int32_t si : a.seq_index) {
      const Sequence& s = seqs_[si];
      const int32_t n = s.len();
      const int32_t start = std::min(s.num_cached_tokens, n - 1);
      any_cached |= start > 0;
      for (int32_t p = start; p < n; ++p) {
        a.input_ids.push_back(s.token_ids[p]);
        a.positions.push_back(p);
        a.slot_mapping.push_back(s.block_table[p / block_size_] * block_size_ + p % block_size_);
      }
      a.cu_q.push_back(a.cu_q.back() + (n - start));
      a.cu_k.push_back(a.cu_k.back() + n);
      max_q = std::max(max_q, n - start);
      max_k = std::max(max_k, n);
    }
    out->num_tokens = static_cast<int32_t>(a.input_ids.size());
    out->input_ids = a.input_ids.data(); out->positions = a.positions.data();
    out->slot_mapping = a.slot_mapping.data();
    out->cu_seqlens_q = a.cu_q.data(); out->cu_seqlens_k = a.cu_k.data();
    out->max_seqlen_q = max_q; out->max_seqlen_k = max_k;
    if (any_cached) fill_tables(out);
  }

  // prepare_decode_inputs + create_decode_context (model_runner.rs:196-210,266-300)
  void build_decode(nvllm_batch* out) {
    fill_common(out, false);
    BatchArrays& a = arrays_;
    const size_t n = a.seq_index.size();
    a.input_ids.resize(n); a.positions.resize(n); a.slot_mapping.resize(n); a.context_lens.resize(n);
    for (size_t i = 0; i < n; ++i) {
      const Sequence& s = seqs_[a.seq_index[i]];
      a.input_ids[i] = s.last_token;                                                          // :200
      a.positions[i] = s.len() - 1;                                                           // :201
      a.slot_mapping[i] = s.block_table.back() * block_size_ + s.last_block_num_tokens() - 1; // intent (a14)
      a.context_lens[i] = s.len();                                                            // :275
    }
    out->num_tokens = static_cast<int32_t>(n);
    out->input_ids = a.input_ids.data(); out->positions = a.positions.data();
    out->slot_mapping = a.slot_mapping.data(); out->context_lens = a.context_lens.data();
    fill_tables(out);
  }

  int32_t max_num_seqs_, max_num_batched_tokens_, max_model_len_;
  int64_t eos_;
  int32_t block_size_;
  BlockManager bm_;
  std::vector<Sequence> seqs_;           // slab; queues hold slab indices, id_to_slot_ maps seq_id -> index
  std::unordered_map<int64_t, int32_t> id_to_slot_;
  std::vector<int32_t> free_slots_;      // slab entries of released sequences, reused by add_sequence
  int64_t next_seq_id_ = 0;
  std::deque<int32_t> waiting_, running_;
  BatchArrays arrays_;
  int64_t total_sequences_ = 0, finished_ = 0, preemptions_ = 0, prefill_batches_ = 0, decode_batches_ = 0;
};

}  // namespace host
}  // namespace nvllm
