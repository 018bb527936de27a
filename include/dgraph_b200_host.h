/* C ABI of libdgraph_b200_host.so -- the host-only runtime (no CUDA, no torch): the multilevel graph partitioner and the
 * one-sided shared-memory windows of the CPU engine.  Sources: csrc/host/partition.cpp, csrc/host/shm_window.cpp.
 * Python users go through dgraph_b200.data.partition.multilevel and dgraph_b200.parallel.shm_engine. */
#ifndef DGRAPH_B200_HOST_H
#define DGRAPH_B200_HOST_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- graph partitioner (the METIS role; reference: experiments/OGB/preprocess.py:15-29) ---------------------------------------
 * n vertices; CSR of the undirected graph (rowptr int64 [n+1], col int32 [rowptr[n]]); vweights int64 [n] or NULL (= 1 each);
 * k parts; imbalance = allowed largest part weight / ideal (e.g. 1.03); refine_iters boundary-refinement sweeps per level;
 * use_triangle_weights: coarsen along common-neighbour-weighted edges; part_out int32 [n].
 * Returns the number of cut edges (>= 0), or -1 when the arguments are rejected. */
int64_t dgb_partition_kway(int64_t n, const int64_t* rowptr, const int32_t* col, const int64_t* vweights, int k, double imbalance,
                           int refine_iters, int use_triangle_weights, uint64_t seed, int32_t* part_out);

/* ---- one-sided windows over POSIX shared memory (the MPI-RMA role; reference: DGraph/distributed/mpi/MPIBackendEngine.py:33-193) -
 * Every call below returns 0 on success, else: 1 timed out waiting for a peer, 2 row outside the owner's window, 3 a system
 * call failed (errno on stderr), 4 bad argument, 5 another rank reported a failure.
 *
 * An access epoch is: [owner prepares its window] fence [gets / puts / accumulates] fence [owner reads its window].  Nothing is
 * allocated or mapped between the two fences. */

/* This rank's control block and first data segment ("/<job>_c<rank>", "/<job>_d<rank>_<generation>"); `threads` copy /
 * accumulate threads for large requests.  NULL on failure.  Does not wait for the peers: dgbh_win_connect does. */
void* dgbh_win_open(const char* job, int rank, int world, uint64_t bytes, int threads);
/* Map every peer's control block and data segment; waits up to timeout_ms for peers that have not created theirs yet. */
int dgbh_win_connect(void* win, double timeout_ms);
void* dgbh_win_base(void* win);     /* this rank's data segment */
uint64_t dgbh_win_bytes(void* win); /* its capacity */
uint64_t dgbh_win_epoch(void* win); /* fences passed */
/* Replace this rank's segment by a larger one (contents are not kept).  Only outside an access epoch; peers re-map it at their
 * next fence. */
int dgbh_win_grow(void* win, uint64_t bytes);
/* Epoch fence: what any rank wrote before its call is visible to every rank after it.  Bounded wait; a timeout marks the whole
 * window set aborted. */
int dgbh_win_fence(void* win, double timeout_ms);
/* dst[dst_row ? dst_row[i] : i] = window(owner[i])[base_off + row[i] * row_bytes ...]   for i < n */
int dgbh_win_get_rows(void* win, const int64_t* row, const int32_t* owner, const int64_t* dst_row, int64_t n, int64_t row_bytes,
                      int64_t base_off, void* dst);
/* window(peer)[dst_off ...] = src[0 .. nbytes) */
int dgbh_win_put(void* win, int peer, int64_t dst_off, const void* src, int64_t nbytes);
/* window(owner[i])[base_off + row[i] * F ...] += src[i * F ...]; elem_bytes 4 (float) or 8 (double).  Rows of an owner that
 * published lock bytes are locked one at a time and added with plain vector adds; otherwise every element is added with a
 * compare-and-swap loop.  Safe against concurrent accumulates from any number of ranks and threads. */
int dgbh_win_accumulate_rows(void* win, const int64_t* row, const int32_t* owner, int64_t n, int64_t F, int elem_bytes,
                             int64_t base_off, const void* src);
/* Publish the byte offset of this rank's per-row lock bytes (one zeroed byte per row) for the coming accumulate epoch; -1
 * withdraws them.  Takes effect for the peers at the next fence. */
int dgbh_win_set_lock_offset(void* win, int64_t off);
int dgbh_win_aborted(void* win); /* 1 once any rank gave up waiting */
void dgbh_win_close(void* win);  /* unmap everything, unlink this rank's segments */

#ifdef __cplusplus
}
#endif
#endif
