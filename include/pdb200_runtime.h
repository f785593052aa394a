/* pdb200_runtime.h — host-side runtime of pytorch_distributed_b200 (no CUDA dependency).
 *
 * C ABI so that Python (ctypes), C and C++ callers share one surface.  Everything here is
 * plain host code: the gradient-bucket planner and reducer state machine that drive the
 * fused all-reduce kernels, and NUMA/CPU-affinity binding for spawned ranks.
 *
 * Parity map:
 *   bucket planner / reducer  <->  torch c10d Reducer (reducer.hpp:45-202,276-336) and
 *                                  _compute_bucket_assignment_by_size
 *   numa binding              <->  hfai.multiprocessing.spawn(bind_numa=True)
 *                                  (ref:restnet_ddp.py:155)
 */
#ifndef PDB200_RUNTIME_H_
#define PDB200_RUNTIME_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- bucket planner ---------------------------------------------------------------
 * Parameters live in one arena; parameter i occupies elements [offset[i], offset[i]+numel[i]).
 * Gradients become ready roughly in reverse definition order, so buckets are contiguous
 * arena ranges assigned walking i = n-1 .. 0: the first bucket closes once it holds
 * >= first_cap_bytes, later ones at >= cap_bytes (torch: 1 MiB / 25 MiB, reducer.hpp:30-31).
 * Writes, in launch order, bucket_start/bucket_end (element ranges, end rounded up to
 * `align` elements) and param_bucket[i].  Returns the number of buckets, or -1 if
 * max_buckets is too small / arguments are invalid. */
int pdb_plan_buckets(int n_params, const int64_t* offset, const int64_t* numel, int elem_bytes,
                     int64_t first_cap_bytes, int64_t cap_bytes, int64_t align, int max_buckets,
                     int64_t* bucket_start, int64_t* bucket_end, int32_t* param_bucket);

/* ---- reducer state machine ----------------------------------------------------------
 * Tracks which parameters have produced their gradient this iteration and releases
 * buckets strictly in launch order (bucket k is released only after 0..k-1), which is what
 * keeps every rank issuing collectives in the same order (reducer.hpp:299-301). */
typedef struct pdb_reducer pdb_reducer;

pdb_reducer* pdb_reducer_create(int n_params, const int32_t* param_bucket, int n_buckets);
void pdb_reducer_destroy(pdb_reducer* r);
/* Start a new backward pass: clears ready marks. */
void pdb_reducer_reset(pdb_reducer* r);
/* Mark parameter `param` ready.  Writes up to max_out released bucket indices to out (in
 * launch order) and returns how many were released; -1 if the parameter was already marked
 * in this pass (gradient produced twice: unsupported without re-arming) or out of range. */
int pdb_reducer_mark_ready(pdb_reducer* r, int param, int32_t* out, int max_out);
/* Release every remaining bucket (end of backward with unused parameters). */
int pdb_reducer_flush(pdb_reducer* r, int32_t* out, int max_out);
/* 1 when all buckets have been released in this pass. */
int pdb_reducer_done(const pdb_reducer* r);
/* Number of parameters still pending in bucket b. */
int pdb_reducer_pending(const pdb_reducer* r, int bucket);

/* ---- NUMA / CPU affinity ------------------------------------------------------------
 * Binds the calling process to the CPUs of the NUMA node that hosts the PCI device
 * `pci_bus_id` (e.g. "0000:1b:00.0", as returned by cudaDeviceGetPCIBusId) by reading sysfs.
 * Returns the node (>= 0), -1 if the node is unknown (single-socket / virtualised: nothing is
 * changed), -2 on sched_setaffinity failure. */
int pdb_bind_numa_for_pci(const char* pci_bus_id);
/* Same, given the node directly. */
int pdb_bind_numa_node(int node);
/* Number of CPUs in the calling thread's affinity mask. */
int pdb_affinity_count(void);

/* Library version, for load-time sanity checks. */
int pdb_runtime_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PDB200_RUNTIME_H_ */
