// Gradient bucket planner + reducer state machine (host side).
//
// Counterpart of torch's c10d Reducer bookkeeping (reducer.hpp:45-202, 276-336) — but only the
// bookkeeping: because gradients already live in an arena, a "bucket" is an address range and
// there is nothing to flatten or copy; the reducer only decides *when* a range may be handed
// to the fused all-reduce kernel, and keeps the launch order identical on every rank.
#include "pdb200_runtime.h"

#include <cstdlib>
#include <mutex>
#include <vector>

extern "C" int pdb_runtime_version(void) { return 1; }

extern "C" int pdb_plan_buckets(int n_params, const int64_t* offset, const int64_t* numel,
                                int elem_bytes, int64_t first_cap_bytes, int64_t cap_bytes,
                                int64_t align, int max_buckets, int64_t* bucket_start,
                                int64_t* bucket_end, int32_t* param_bucket) {
  if (n_params <= 0 || !offset || !numel || elem_bytes <= 0 || align <= 0 || max_buckets <= 0)
    return -1;
  int nb = 0;
  bool open = false;
  int64_t cur_end = 0, cur_start = 0, cap = first_cap_bytes;
  for (int i = n_params - 1; i >= 0; --i) {
    if (i + 1 < n_params && offset[i] + numel[i] > offset[i + 1]) return -1;  // must be ordered
    const int64_t end = (offset[i] + numel[i] + align - 1) / align * align;
    if (!open) {
      if (nb >= max_buckets) return -1;
      cur_end = end;
      open = true;
    }
    cur_start = offset[i];
    param_bucket[i] = nb;
    if ((cur_end - cur_start) * elem_bytes >= cap) {
      bucket_start[nb] = cur_start;
      bucket_end[nb] = cur_end;
      ++nb;
      open = false;
      cap = cap_bytes;
    }
  }
  if (open) {
    bucket_start[nb] = cur_start;
    bucket_end[nb] = cur_end;
    ++nb;
  }
  return nb;
}

struct pdb_reducer {
  std::vector<int32_t> param_bucket;
  std::vector<int32_t> bucket_size;     // parameters per bucket
  std::vector<int32_t> pending;         // still-missing parameters per bucket, this pass
  std::vector<uint8_t> ready;           // per parameter, this pass
  int next_bucket = 0;                  // first bucket not yet released
  std::mutex mu;                        // autograd may call hooks from several threads (DP mode)
};

extern "C" pdb_reducer* pdb_reducer_create(int n_params, const int32_t* param_bucket, int n_buckets) {
  if (n_params <= 0 || n_buckets <= 0 || !param_bucket) return nullptr;
  auto* r = new pdb_reducer();
  r->param_bucket.assign(param_bucket, param_bucket + n_params);
  r->bucket_size.assign(n_buckets, 0);
  for (int i = 0; i < n_params; ++i) {
    if (param_bucket[i] < 0 || param_bucket[i] >= n_buckets) {
      delete r;
      return nullptr;
    }
    r->bucket_size[param_bucket[i]]++;
  }
  r->ready.assign(n_params, 0);
  r->pending = r->bucket_size;
  return r;
}

extern "C" void pdb_reducer_destroy(pdb_reducer* r) { delete r; }

extern "C" void pdb_reducer_reset(pdb_reducer* r) {
  std::lock_guard<std::mutex> g(r->mu);
  r->pending = r->bucket_size;
  std::fill(r->ready.begin(), r->ready.end(), 0);
  r->next_bucket = 0;
}

static int release_locked(pdb_reducer* r, int32_t* out, int max_out, bool force) {
  int n = 0;
  const int nb = static_cast<int>(r->pending.size());
  while (r->next_bucket < nb && (force || r->pending[r->next_bucket] == 0)) {
    if (n < max_out && out) out[n] = r->next_bucket;
    ++n;
    ++r->next_bucket;
  }
  return n;
}

extern "C" int pdb_reducer_mark_ready(pdb_reducer* r, int param, int32_t* out, int max_out) {
  std::lock_guard<std::mutex> g(r->mu);
  if (param < 0 || param >= static_cast<int>(r->ready.size()) || r->ready[param]) return -1;
  r->ready[param] = 1;
  r->pending[r->param_bucket[param]]--;
  return release_locked(r, out, max_out, false);
}

extern "C" int pdb_reducer_flush(pdb_reducer* r, int32_t* out, int max_out) {
  std::lock_guard<std::mutex> g(r->mu);
  return release_locked(r, out, max_out, true);
}

extern "C" int pdb_reducer_done(const pdb_reducer* r) {
  return r->next_bucket >= static_cast<int>(r->pending.size()) ? 1 : 0;
}

extern "C" int pdb_reducer_pending(const pdb_reducer* r, int bucket) {
  if (bucket < 0 || bucket >= static_cast<int>(r->pending.size())) return -1;
  return r->pending[bucket];
}
