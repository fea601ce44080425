/* bqa_b200.h - C ABI of the B200-native emulator backend for bloqade-analog's Python-emulator hot path.
 *
 * The reference (QuEraComputing/bloqade-analog) has NO FFI for this path: it is pure Python that
 * reaches native code only through SciPy / Numba.  This header is therefore the
 * boundary a maintainer would bind with ctypes from the reference's own classes; each entry point
 * names the reference function(s) it replaces (paths relative to src/bloqade/analog/).
 * INTEGRATION.md shows the reference-side stub.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on error; bqa_last_error() gives the message
 *     (thread-local).  No exceptions cross the boundary.
 *   - state vectors are complex128, interleaved (re, im), contiguous - numpy complex128 layout.
 *   - `mem` says where the DATA buffers of a call live: BQA_MEM_HOST (the library stages them
 *     through the device: H2D/D2H inside the call) or BQA_MEM_DEVICE (used in place, no copies).
 *     Plan descriptors, solver options, stats and scalar results are always host memory.
 *   - work is enqueued on the plan's stream (bqa_plan_set_stream, default stream 0) and the call
 *     returns after the stream is synchronised when any output is host memory.
 *   - a plan belongs to the CUDA device that was current when it was created; every entry point that
 *     takes a plan makes that device current in the calling thread, so one host thread per GPU can drive
 *     plans of different devices concurrently (the routine's `devices=` sharding does).
 */
#ifndef BQA_B200_H
#define BQA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BQA_MEM_HOST 1
#define BQA_MEM_DEVICE 2

#define BQA_SOLVER_DOPRI5 5
#define BQA_SOLVER_DOP853 8

/* integrator status per task = scipy's dopri5/dop853 return code (emulate/ir/state_vector.py:494-508) */
#define BQA_STATUS_OK 1
#define BQA_STATUS_INCONSISTENT (-1)
#define BQA_STATUS_NEED_STEPS (-2)
#define BQA_STATUS_STEP_TOO_SMALL (-3)
#define BQA_STATUS_STIFF (-4)
#define BQA_STATUS_PEER_ABORT (-6) /* sharded state only: a peer rank failed or a cross-rank barrier timed out */

#define BQA_ENGINE_AUTO 0
#define BQA_ENGINE_CTA_PER_TASK 1 /* kernel (5): whole integration inside one CTA per task        */
#define BQA_ENGINE_STREAMED 2     /* kernels (1)+(2): multi-CTA stage kernels, device controller  */
#define BQA_ENGINE_CLUSTER 3      /* kernel (5): one thread-block cluster per task, all vectors in
                                     distributed shared memory (AUTO picks it whenever it fits)    */

typedef struct bqa_plan bqa_plan;

/* Hilbert space of a register.  Replaces the arrays held by emulate/ir/space.py:21-27 (Space) and the
 * index maps of Space.swap_state_at / transition_state_at (space.py:134-178) in gather form.
 * Full spaces (configuration == index) pass configs = partners = NULL. */
typedef struct {
  int32_t n_atoms;
  int32_t n_level;          /* 2 (g,r) or 3 (g,h,r); atom 0 = least significant digit              */
  int64_t dim;
  const uint32_t* configs;  /* [dim] sorted configurations of a blockade subspace, or NULL          */
  const int32_t* partners;  /* [n_atoms][n_level-1][dim]: index of the state with digit s<->s+1
                               exchanged on atom a, -1 if none; or NULL                             */
} bqa_space_desc;

typedef struct {
  int32_t solver;              /* BQA_SOLVER_*  (AnalogGate.SUPPORTED_SOLVERS, state_vector.py:489)  */
  int32_t interaction_picture; /* state_vector.py:604-641                                            */
  double atol, rtol;
  int64_t nsteps;
  int32_t reject_mode;         /* DOP853 step after a rejection: 1 = SciPy >= 1.17 C translation (the
                                  Python-side default: same step sequence and failure codes as the
                                  oracle), 0 = Hairer dop853.f / SciPy <= 1.16 (fewer steps)         */
  int32_t engine;              /* BQA_ENGINE_*                                                       */
  double t_start;              /* time of psi0 (0 in the reference's flow).  Non-zero: resume of a run that was
                                  checkpointed at an output time - SciPy's ode restarts the integrator at every
                                  output time (state_vector.py:589-600), so evolving to t1, saving the state and
                                  resuming with t_start = t1 continues bit for bit like the uninterrupted run    */
} bqa_solver_opts;

typedef struct {
  int32_t status;
  int32_t pad;
  int64_t nfcn, nstep, naccpt, nrejct;
  double h_last, t_last;
} bqa_task_stats;

/* ---- library / device ------------------------------------------------------------------------- */
int bqa_version(void);
const char* bqa_last_error(void);
int bqa_device_count(int32_t* count);
int bqa_set_device(int32_t device);
int bqa_current_device(int32_t* device);
int bqa_device_info(int32_t* sm_count, int64_t* smem_per_block_optin, int64_t* l2_bytes);
/* total kernel launches issued by this library since load (bench.py "gpu_launches") */
int64_t bqa_launch_count(void);
/* Pass list of the shared-memory-tiled H(t) psi (csrc/bqa_apply_tiled.cuh) for the n_local atoms a device holds
 * of an n_total-atom two-level register (n_total > n_local: sharded state, rank 0).  Host arithmetic only - no
 * GPU needed; the CPU test suite checks with it that every atom is flipped exactly once and every tile is
 * visited exactly once.  passes[i][8] = {lo_bits, hb, tb, p0, n_x, xbit, pair, pair_b}.  Returns the number
 * of passes (0: this shape is not tiled), or -1 when max_passes is too small. */
int bqa_tiled_plan_describe(int32_t n_local, int32_t n_total, int32_t* passes, int32_t max_passes);

/* ---- plan: structure shared by all tasks of a sweep ---------------------------------------------
 * replaces RydbergHamiltonianCodeGen.visit_register (emulate/codegen/hamiltonian.py:63-91) */
int bqa_plan_create(const bqa_space_desc* space, bqa_plan** plan);
void bqa_plan_destroy(bqa_plan* plan);
int bqa_plan_set_stream(bqa_plan* plan, void* cuda_stream);
/* static (van-der-Waals) diagonal `rydberg` f64[dim], host memory; returns its id (task record field) */
int bqa_plan_add_diagonal(bqa_plan* plan, const double* v_host, int32_t* diag_id);
/* same diagonal built ON DEVICE from pair[a][b] = C6/|r_a-r_b|^6 (a<b, row-major [n][n], 0 = cut),
 * bit-identical to hamiltonian.py:76-89 (same pair order, additions only) */
int bqa_plan_build_rydberg(bqa_plan* plan, const double* pair_host, int32_t* diag_id);
int bqa_plan_get_diagonal(bqa_plan* plan, int32_t diag_id, double* v_host);
/* generic SparseMatrixCSR / CSC fallback term (emulate/sparse_operator.py:15-34,68-96): M in CSR with
 * f64 data; applied as A(t) M psi and, when phase_wf >= 0, + conj(A(t)) M^T psi (state_vector.py:279-288).
 * amp_wf / phase_wf index the waveforms of the task record. */
int bqa_plan_add_csr(bqa_plan* plan, int32_t amp_wf, int32_t phase_wf, int64_t nnz, const double* data,
                     const int32_t* indices, const int64_t* indptr);

/* instrumentation for bench.py: when enabled, the evolve engines bracket their kernels with CUDA events
 * recorded on the plan's stream; get_timing returns the cumulative device time (ms) and the number of
 * bracketed kernel launches since timing was (re-)enabled. */
int bqa_plan_enable_timing(bqa_plan* plan, int32_t on);
int bqa_plan_get_timing(bqa_plan* plan, double* evolve_kernel_ms, int64_t* evolve_kernel_launches);

/* ---- task record layout (float64 array; integers stored as exact doubles) -----------------------
 * replaces the DetuningOperator / RabiOperator lists (state_vector.py:261-298) and the JITWaveform
 * callables (emulate/ir/emulator.py:65-79):
 *   [0] duration  [1] diag_id (-1 none)  [2] n_wf  [3] n_pieces  [4] n_det  [5] n_rabi  [6] n_coef  [7] n_atoms
 *   wf_start[n_wf+1]                      first piece of each waveform
 *   pieces[n_pieces][6]                   lo, hi, offset, scale, coef_start, n_coef:
 *                                         value(t) = scale * poly(t - offset) if lo < t <= hi else 0
 *   coefs[n_coef]                         polynomial coefficients, lowest order first
 *   det[n_det][2 + n_atoms]               wf, slot, scale per atom   (diagonal -= f(t)*scale on digit slot+1)
 *   rabi[n_rabi][3 + n_atoms]             amp_wf, phase_wf (-1 none), slot, scale per atom
 *                                         (exchange digits slot <-> slot+1 with (Omega/2) e^{-i phi} * scale)
 * A batch is the concatenation of records plus offsets[n_tasks+1] (in doubles). */

/* ---- kernel (1): one fused H(t) psi ----------------------------------------------------------------
 * replaces RydbergHamiltonian._ode_complex_kernel / _ode_complex_kernel_int (state_vector.py:323-359)
 * when raw == 0 (out = -i H psi) and RydbergHamiltonian._apply (state_vector.py:375-398) when raw != 0.
 * `record` is always HOST memory (it is small); psi_in / psi_out follow `mem`. */
int bqa_apply(bqa_plan* plan, const double* record, int64_t record_len, double t, int32_t interaction_picture,
              int32_t raw, const void* psi_in, void* psi_out, int32_t mem);

/* ---- kernels (2)+(5): adaptive evolution of a batch of tasks -----------------------------------------
 * replaces AnalogGate._apply / _apply_interaction_picture (state_vector.py:574-641) and the task loop of
 * LocalBatch._run (task/batch.py:251-252).  times[n_tasks][n_times]; psi_out[n_tasks][n_times][dim];
 * psi0: NULL = |0..0> (Space.zero_state, space.py:204-209); psi0_shared != 0: one psi0[dim] for all tasks.
 * records / offsets / times / psi0 / psi_out follow `mem`; stats[n_tasks][n_times] is host memory. */
int bqa_evolve_batch(bqa_plan* plan, const double* records, const int64_t* offsets, int32_t n_tasks,
                     const bqa_solver_opts* opts, const void* psi0, int32_t psi0_shared, const double* times,
                     int32_t n_times, void* psi_out, bqa_task_stats* stats, int32_t mem);

/* ---- kernels (3)+(4) fused behind AnalogGate.run (state_vector.py:674-699): evolve every task to its
 * `times` entry, normalise, Born-sample n_shots per task with the given uniforms[n_tasks][n_shots]
 * (np.random.random_sample - what np.random.choice draws, space.py:217) -> shots[n_tasks][n_shots][n_atoms]
 * uint8 digits (project_hyperfine: h->0, r->1).  psi_out (optional, may be NULL) receives the normalised
 * final states. */
int bqa_run_batch(bqa_plan* plan, const double* records, const int64_t* offsets, int32_t n_tasks,
                  const bqa_solver_opts* opts, const double* times, const double* uniforms, int32_t n_shots,
                  int32_t project_hyperfine, uint8_t* shots, void* psi_out, bqa_task_stats* stats, int32_t mem);

/* ---- kernel (3): reductions ------------------------------------------------------------------------ */
/* StateVector.norm (state_vector.py:215-217) */
int bqa_norm(bqa_plan* plan, const void* psi, int32_t mem, double* norm);
/* StateVector.normalize, in place (state_vector.py:210-213) */
int bqa_normalize(bqa_plan* plan, void* psi, int32_t mem);
/* RydbergHamiltonian.average_and_variance (state_vector.py:419-442): avg = Re<psi|H psi>,
 * var = <H psi|H psi> - avg^2 */
int bqa_expectation(bqa_plan* plan, const double* record, int64_t record_len, double t, const void* psi,
                    int32_t mem, double* average, double* variance);
/* StateVector.local_trace (state_vector.py:88-175): n_sites 1 or 2; op is the dense complex
 * (n_level^n_sites)^2 matrix, row-major, host memory; sites are ATOM indices (already mapped through
 * full_index_to_index).  result[2] = (re, im) of the raw sum (the 1-body division by norm() is host side). */
int bqa_local_trace(bqa_plan* plan, const void* psi, int32_t mem, int32_t n_sites, int32_t site_a, int32_t site_b,
                    const double* op, double* result);

/* ---- allocation-free norm ------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same two
 * deterministic passes as bqa_norm (StateVector.norm), but the per-CTA partial sums live in a file-scope
 * `static __device__` array of BQA_NOALLOC_MAX_PARTIALS doubles (512 KiB, held by every GPU that loads the
 * library until the process exits) instead of cudaMalloc'ed plan scratch.  The call allocates nothing, copies
 * nothing to the host and does not synchronise: psi_dev[n_states][dim] and norm_dev[n_states] are DEVICE
 * pointers owned by the caller, the two kernels are enqueued on the plan's stream and norm_dev[s] = ||psi_s||_2
 * is valid once that stream reaches the end of the call's work.  The plan's device must be current.  Rejected
 * with an error when n_states * (CTAs per state, at most 4 per SM) exceeds BQA_NOALLOC_MAX_PARTIALS.  All
 * calls on one GPU share the array: calls that overlap on the GPU (different streams / threads) race; calls
 * ordered on one stream do not.  Nothing has to start at zero (every slot read was written by the same call).
 * The first call on a device looks the array's address up (cudaGetSymbolAddress): make it outside a capture. */
#define BQA_NOALLOC_MAX_PARTIALS 65536
int bqa_norm_noalloc(bqa_plan* plan, const void* psi_dev, int32_t n_states, double* norm_dev);

/* ---- kernel (4): Born sampling (Space.sample_state_vector, space.py:211-229) ---------------------------
 * psi[n_states][dim] (normalised or not: the CDF is divided by its last element like np.random.choice),
 * uniforms[n_states][n_shots] -> shots[n_states][n_shots][n_atoms] */
int bqa_sample(bqa_plan* plan, const void* psi, int32_t n_states, const double* uniforms, int32_t n_shots,
               int32_t project_hyperfine, uint8_t* shots, int32_t mem);

/* ---- sharded full-space state (BASELINE config 5) ------------------------------
 * No reference counterpart: the reference keeps the state in ONE NumPy array (emulate/ir/state_vector.py:88-90)
 * and cannot hold 2^30 amplitudes with its CSR operators (emulate/codegen/hamiltonian.py:161).  Rank r of
 * `world` = 2^g ranks (one process per GPU) owns the amplitudes whose top g index bits equal r (atom k <-> index
 * bit k, atom_type.py:139,151); the plan of every rank describes the n_atoms_total - g LOCAL atoms, the task
 * record all n_atoms_total atoms.  Flips of a rank-bit atom read the partner shard in place over NVLink peer
 * access; reductions and barriers of the integrator are device initiated (no host round trip, no collective
 * library on the data path).  Buffers other ranks must see are allocated with bqa_shard_alloc (cudaMalloc +
 * CUDA IPC handle, zero filled); the 64-byte handles are exchanged by the host (torch.distributed
 * all_gather_object) and opened with bqa_shard_open. */
typedef struct {
  int32_t world, rank;      /* world = 2, 4 or 8                                                          */
  int32_t n_atoms_total;    /* atoms of the whole register                                                */
  int32_t pad;
  void* ys[8];              /* [rank]: that rank's stage-input buffer, local_dim complex128 (own or peer-mapped) */
  void* sync[8];            /* [rank]: that rank's sync block, bqa_shard_sync_bytes() bytes                */
} bqa_shard_desc;

int64_t bqa_shard_sync_bytes(void);
int bqa_shard_alloc(int64_t bytes, void** dev_ptr, void* handle64);
int bqa_shard_open(const void* handle64, void** peer_ptr);
int bqa_shard_close(void* peer_ptr);
int bqa_shard_free(void* dev_ptr);
/* Zero the barrier counters and the abort word of THIS rank's sync block.  Every rank calls it, then a host-side
 * barrier (torch.distributed / MPI), before each bqa_evolve_sharded: the cross-rank barriers of the integrator
 * are device initiated with a bounded spin (BQA_BARRIER_TIMEOUT_MS, default 60000) and an abort word, so a rank
 * that fails or stalls ends the others with status BQA_STATUS_PEER_ABORT (-6) instead of hanging them, and the
 * reset brings the ranks back in step afterwards.  No reference counterpart (single-process NumPy state). */
int bqa_shard_reset(void* sync_own);
/* van-der-Waals diagonal of this rank's shard, same pair order as hamiltonian.py:76-89; pair_host is
 * [n_atoms_total][n_atoms_total] */
int bqa_plan_build_rydberg_shard(bqa_plan* plan, int32_t n_atoms_total, const double* pair_host, int32_t rank,
                                 int32_t* diag_id);
/* out_local = [-i] (H(t) psi)_local; partners[b] = device pointer of the shard of rank ^ (1 << b); the caller
 * guarantees those shards are complete (no barrier inside).  Device pointers only. */
int bqa_apply_sharded(bqa_plan* plan, const double* record, int64_t record_len, double t, int32_t raw,
                      int32_t n_atoms_total, int32_t rank, const void* const* partners, const void* psi_in,
                      void* psi_out);
/* AnalogGate._apply (state_vector.py:574-602) for one sharded state; every rank calls it with the same record,
 * options and times.  psi0_local: device pointer of this rank's shard, or NULL for |0...0>; psi_out_local:
 * device, [n_times][local_dim]; stats: host, [n_times]. */
int bqa_evolve_sharded(bqa_plan* plan, const double* record, int64_t record_len, const bqa_solver_opts* opts,
                       const bqa_shard_desc* shard, const void* psi0_local, const double* times, int32_t n_times,
                       void* psi_out_local, bqa_task_stats* stats);

#ifdef __cplusplus
}
#endif
#endif /* BQA_B200_H */
