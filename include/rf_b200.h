/* rf_b200.h -- C ABI of librf_b200.so, the B200-native field-synthesis path.
 *
 * The reference (dkirkby/randomfield) is pure Python and has no FFI of its own; its
 * de-facto operator interface for this path is the set of in-place array functions that
 * Generator.generate_delta_field calls (randomfield/generate.py:191-219) plus
 * transform.Plan (randomfield/transform.py:161-315).  Each entry point below names the
 * reference interface it replaces.  All functions return 0 on success and a negative
 * status otherwise; rf_last_error() returns a human-readable message for the calling
 * thread.  Plain pointers and sizes only -- no C++ or torch types.
 *
 * A context owns the device buffers of one Generator (one rank of a slab-decomposed grid).
 * Calls on one context must be serialised by the caller; work is asynchronous on the
 * context's CUDA stream until rf_sync / a host copy.
 */
#ifndef RF_B200_H
#define RF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rf_ctx rf_ctx;

enum { RF_F32 = 0, RF_F64 = 1 };
enum { RF_OK = 0, RF_ERR_ARG = -1, RF_ERR_CUDA = -2, RF_ERR_STATE = -3 };

/* which device buffer rf_device_ptr returns.  RF_BUF_AUX is the same memory as RF_BUF_RECV: the
 * all-to-all receive buffer during a slab-decomposed step, afterwards the second real-space array
 * (the lensing potential of rf_lensing; scratch of rf_r2c_power) */
enum { RF_BUF_FIELD = 0, RF_BUF_RECV = 1, RF_BUF_PARTIALS = 2, RF_BUF_AUX = 1 };

const char* rf_last_error(void);
int rf_version(void);

/* transform.Plan.__init__ / Generator.__init__ buffer allocation (transform.py:170-244,
 * generate.py:76-80).  Grid (nx, ny, nz), all even.  rank/world describe the slab
 * decomposition: this rank generates the k-space y-slab [rank*ny/world, (rank+1)*ny/world)
 * and ends with the real-space x-slab [rank*nx/world, ...).  world == 1 for a single GPU. */
int rf_create(int nx, int ny, int nz, int dtype, int rank, int world, rf_ctx** out);
int rf_destroy(rf_ctx* ctx);

/* run on an externally owned CUDA stream (cudaStream_t passed as void*); NULL restores
 * the context's own stream */
int rf_set_stream(rf_ctx* ctx, void* cuda_stream);
int rf_sync(rf_ctx* ctx);

/* powertools.tabulate_sigmas table (powertools.py:152-157): n knots, k[] strictly
 * increasing and > 0, sigma[] = N3*sqrt(Pk/(2*Vbox)) in the reference normalisation
 * (smoothing from powertools.filter_power already applied by the caller). */
int rf_set_power(rf_ctx* ctx, int n, const double* k, const double* sigma, double spacing);

/* White-noise injection (replaces the RandomState(seed).normal draw of random.py:24-28):
 * host array of 2*nx*ny*(nz/2+1) doubles in the reference order, uploaded once.
 * NULL clears it and returns to the Philox generator. */
int rf_set_noise(rf_ctx* ctx, const double* host_noise);
/* the same array uploaded in slabs of x rows [x0, x1) (host_rows points at row x0), so that the
 * caller never has to hold 16 bytes per mode on the host (random.py:25-27: "should chunk") */
int rf_set_noise_rows(rf_ctx* ctx, const double* host_rows, int x0, int x1);

/* Pass 1 of generate_delta_field: fill_with_log10k + tabulate_sigmas + randomize +
 * symmetrize (generate.py:191-199) fused with the axis-0 pass of Plan.execute
 * (transform.py:303-315).  potential != 0 scales every mode by 1/k^2 (generate.py:200-215).
 * Output: this rank's [nx][ny/world][nz/2] block in RF_BUF_FIELD. */
int rf_generate_xpass(rf_ctx* ctx, uint64_t seed, int potential);

/* Passes 2+3 (axis 1, then c2r along axis 2).  For world > 1 the caller first exchanges
 * RF_BUF_FIELD -> RF_BUF_RECV with an all-to-all of world equal blocks.  zscale (nz host
 * doubles or NULL) multiplies the output per z index (generate.py:341-343); want_stats
 * accumulates sum / sum-of-squares of the unscaled field (generate.py:219). */
int rf_finish_yz(rf_ctx* ctx, const double* zscale, int want_stats);

/* single-GPU convenience: rf_generate_xpass + rf_finish_yz */
int rf_generate_delta(rf_ctx* ctx, uint64_t seed, int potential, const double* zscale,
                      int want_stats);

/* np.std(delta.flat) ingredients (generate.py:219): local sum, sum of squares, count.
 * Synchronises the stream. */
int rf_stats(rf_ctx* ctx, double* sum, double* sumsq, long long* count);

/* convert_delta_to_density (generate.py:266-273) incl. apply_lognormal_transform
 * (cosmotools.py:214-221), in place on the real field; growth / rho: nz host doubles. */
int rf_density(rf_ctx* ctx, int lognormal, double sigma, const double* growth,
               const double* rho);

/* copy this rank's real field [nx/world][ny][nz] to host memory (synchronous) */
int rf_copy_field(rf_ctx* ctx, void* host_dst);
/* Streamed output (the reference hands back views of its one work buffer, generate.py:186-189,
 * 224-230; here the result has to cross PCIe).  rf_copy_rows copies local x rows [x0, x1) of
 * RF_BUF_FIELD (the real field) or RF_BUF_AUX (the lensing potential) to host_dst (= address of
 * row x0) on the context's copy stream, ordered after the work queued so far; async != 0 returns
 * at once, rf_copy_wait blocks until every queued copy has landed.  host_dst should be page-locked
 * (rf_host_alloc).  rf_finish_yz_streamed = rf_finish_yz with the output cut into `nranges` ranges
 * of x rows: pass 2 + pass 3 of range r+1 run while range r crosses PCIe (call rf_copy_wait before
 * reading host_dst).  Same kernels on the same data: the field is bit-identical to rf_finish_yz. */
int rf_copy_rows(rf_ctx* ctx, int which, int x0, int x1, void* host_dst, int async);
int rf_copy_wait(rf_ctx* ctx);
int rf_finish_yz_streamed(rf_ctx* ctx, const double* zscale, int want_stats, void* host_dst, int nranges);
/* element count of this rank's real field */
long long rf_field_elems(rf_ctx* ctx);

/* Stage-level access for tests and for the Generator.potential / plan_c2r.data_in
 * attributes: the symmetrised k-space field in the public (nx, ny, nz/2+1) layout and the
 * reference normalisation, copied to host.  bins (optional) receives the canonical
 * sigma-table bin index of every mode. */
int rf_kspace_public(rf_ctx* ctx, uint64_t seed, int potential, void* host_cplx,
                     int32_t* host_bins);

/* transform.Plan.execute (transform.py:303-315) on host arrays: inverse/forward,
 * packed (c2r / r2c with the nz/2+1 layout) or unpacked (c2c), numpy normalisation.
 * in/out are dense C-order host arrays:
 *   inverse packed : in complex (nx,ny,nz/2+1) -> out real (nx,ny,nz)
 *   forward packed : in real (nx,ny,nz)        -> out complex (nx,ny,nz/2+1)
 *   unpacked       : complex (nx,ny,nz) -> complex (nx,ny,nz) */
int rf_plan_execute(int nx, int ny, int nz, int dtype, int inverse, int packed,
                    const void* host_in, void* host_out);

/* stage functions on host arrays in the public layout (ncols = nz/2+1 if packed else nz):
 * powertools.fill_with_log10k (powertools.py:40-61), powertools.tabulate_sigmas
 * (powertools.py:155-163; table = log10 k, sigma), random.randomize (random.py:12-29,
 * Philox instead of MT19937), transform.symmetrize (transform.py:114-158). */
int rf_stage_fill_log10k(int nx, int ny, int nz, int packed, int dtype, double spacing,
                         void* host_cplx);
int rf_stage_tabulate(long long nelem, int dtype, int n, const double* log10k,
                      const double* sigma, void* host_cplx);
int rf_stage_randomize(long long nelem, int dtype, uint64_t seed, void* host_cplx);
int rf_stage_symmetrize(int nx, int ny, int nz, int packed, int dtype, void* host_cplx);

/* cosmotools.apply_lognormal_transform (cosmotools.py:206-221) / the linear branch of
 * convert_delta_to_density (generate.py:271-273) on a host real array whose last axis has
 * nz entries; sigma <= 0 means "use np.std of the data" (computed on the device). */
int rf_stage_density(long long nelem, int nz, int dtype, int lognormal, double sigma,
                     const double* growth, const double* rho, void* host_real);

/* The same epilogue on DEVICE arrays, for callers that cannot allocate (CUDA graph capture,
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * dev_delta (in) and dev_out (out; may be the same array)
 * are caller-owned device arrays of nelem reals, last axis nz; growth / rho are host tables of
 * nz doubles (copied on the stream before the kernels; reusable when the call returns);
 * sigma <= 0 with lognormal != 0 measures np.std(delta) on the device.  The call enqueues
 * two copies and two or three kernels on cuda_stream and returns without synchronising; it
 * calls no cudaMalloc / cudaFree and needs no rf_ctx.  Its scratch (two tables of
 * RF_NOALLOC_MAX_NZ doubles + 1024 partial sums, 80 KB) is a file-scope __device__ array that
 * every GPU loading the library holds until exit: nz > RF_NOALLOC_MAX_NZ is rejected with
 * RF_ERR_ARG.  Every call on a device shares that array and it keeps the last call's contents
 * (each call overwrites all it reads, so nothing needs clearing): calls that overlap on one
 * GPU -- different streams, or host threads without ordering -- race; serialise them. */
#define RF_NOALLOC_MAX_NZ 4096
int rf_density_noalloc(const void* dev_delta, void* dev_out, long long nelem, int nz, int dtype,
                       int lognormal, double sigma, const double* growth, const double* rho,
                       void* cuda_stream);

/* page-locked host memory for result arrays (transform.allocate, transform.py:10-43, uses
 * SIMD-aligned memory for the same reason: fast transfers into the transform) */
void* rf_host_alloc(long long bytes);
void rf_host_free(void* p);

/* per-kernel device timing: when enabled, CUDA events bracket the three hot-path kernels on
 * the context's stream; rf_get_timing synchronises and returns the last step's durations in
 * milliseconds (pass 1 generate+x, pass 2 y, pass 3 z/c2r). */
int rf_set_timing(rf_ctx* ctx, int enabled);
int rf_get_timing(rf_ctx* ctx, double* ms_pass1, double* ms_pass2, double* ms_pass3);

/* Direct exchange (slab decomposition, 2..8 ranks of one box): pass 1 stores every x line
 * straight into the receive buffers of the ranks that own each x range, over NVLink peer
 * memory, so the all-to-all is part of rf_generate_xpass (no send buffer, no separate
 * collective).  Set every rank's RF_BUF_RECV pointer with rf_set_peer_recv (same process: the
 * pointer itself; other processes: rf_ipc_export on the owner, rf_ipc_open here), then
 * rf_set_direct(ctx, 1).  The caller still synchronises the ranks before and after pass 1
 * (peers must have finished reading / writing the receive buffers). */
int rf_ipc_export(rf_ctx* ctx, int which, unsigned char* handle64);
int rf_ipc_open(const unsigned char* handle64, void** ptr);
int rf_ipc_close(void* ptr);
int rf_set_peer_recv(rf_ctx* ctx, int rank, void* recv_ptr);
int rf_set_direct(rf_ctx* ctx, int enabled);

/* Lensing potential psi(r) -- Generator.calculate_lensing_potential (generate.py:353-418):
 * for every slice i, scipy.integrate.simps over DC of -2 (cotK_j - cotK_{i-1}) dPhi_j,
 * j in [i_min, i); slices below i_min are zeroed.  In place on the real-space field (normally the
 * light-cone Newtonian potential), one O(nz) parity-split prefix sum per line of sight.
 * tables5 = 5 contiguous float64 arrays of nz entries built by the caller from DC and cotK:
 *   cotK[j]; a[j], b[j], d[j] = Simpson weights of the three-point panel (j-2, j-1, j) on the
 *   abscissae DC (0 for j < 2); t[j] = (DC[j] - DC[j-1]) / 2 (0 for j = 0).
 * (randomfield_b200.generate.lensing_tables restates the weights of scipy.integrate.simps.)
 * As in the reference (generate.py:398-399) psi is a separate array: it is written to RF_BUF_AUX
 * and the real-space field stays intact (read it back with rf_copy_rows(ctx, RF_BUF_AUX, ...)). */
int rf_lensing(rf_ctx* ctx, int i_min, const double* tables5);

/* Power-spectrum estimator: the forward r2c transform the reference plans but never uses
 * (plan_r2c: generate.py:79-80, transform.py:278-301) of the real-space field, and the mean of
 * |delta(k)|^2 V / N^6 over the modes of the full k-space whose |k| falls in [k_edges[b],
 * k_edges[b+1]) (nbins + 1 edges in h/Mpc).  out_counts (optional) receives the number of modes
 * per shell.  The field is not modified (RF_BUF_AUX is the scratch).  Single-GPU contexts. */
int rf_r2c_power(rf_ctx* ctx, int nbins, const double* k_edges, double* out_pk, long long* out_counts);

/* Chunked exchange (slab decomposition): the packed z columns (nz/2) are cut into `nchunks`
 * chunks.  Pass 1 of chunk c writes a contiguous send block, the copy engines push its G
 * blocks into the peers' receive buffers (rf_set_peer_recv pointers) on the context's comm
 * stream while pass 1 of chunk c+1 and pass 2 of chunk c-1 run on the SMs -- the overlapped
 * form of the global transpose of transform.Plan.execute (transform.py:303-315) when the
 * field is slab-decomposed.  Call order per step (every rank):
 *   rf_chunk_begin_step; [cross-rank fence on rf_comm_stream]
 *   for c: rf_chunk_xpass(c); rf_chunk_exchange(c); [fence on rf_comm_stream]; rf_chunk_arrived(c);
 *          rf_chunk_pass2(c-1)            (pass 2 of a chunk may be queued any time after arrived)
 *   rf_chunk_pass2(last); rf_finish_yz    (pass 3 only)
 * The fences (any cross-rank barrier ordered on the comm stream, e.g. a 1-element NCCL
 * all-reduce) are the caller's: the library has no inter-process transport of its own.
 * world == 1 runs the same kernels without copies (used by the tests). */
int rf_chunk_config(rf_ctx* ctx, int nchunks);          /* 0: off, frees the send buffer */
void* rf_comm_stream(rf_ctx* ctx);
int rf_chunk_begin_step(rf_ctx* ctx);
int rf_chunk_xpass(rf_ctx* ctx, uint64_t seed, int potential, int chunk);
int rf_chunk_exchange(rf_ctx* ctx, int chunk);
int rf_chunk_arrived(rf_ctx* ctx, int chunk);
int rf_chunk_pass2(rf_ctx* ctx, int chunk);
/* Split schedule for transfer-bound decompositions (world >= 4): pass 3 needs every chunk of
 * a line, so with the plan above it cannot start before the LAST chunk has crossed NVLink.
 * Here the exchange is also cut along the destination's local x axis: while pass 1 runs, only
 * rows x in [0, xa) of every chunk are pushed (and pass 2 + pass 3 of those rows follow as soon
 * as pass 1 is over); the remaining rows travel afterwards in x ranges [x0, x1), all chunks of
 * a range in one pitched copy per peer, and pass 2 + pass 3 of one range overlap the transfer of
 * the next.  Call order per step (every rank):
 *   rf_chunk_begin_step; [fence]
 *   for c: rf_chunk_xpass(c); rf_chunk_exchange_x(c, 0, xa, 1); [fence]; rf_chunk_arrived(c);
 *          rf_chunk_pass2_x(c-1, 0, xa, c-1)
 *   rf_chunk_pass2_x(last, 0, xa, last)
 *   for every range s: rf_chunk_exchange_x(-1, x0, x1, 0); [fence]; rf_chunk_arrived(nchunks + s)
 *   rf_chunk_pass3_x(0, xa, ..., last = 0)
 *   for every range s: rf_chunk_pass2_x(-1, x0, x1, nchunks + s);     (every chunk, one launch)
 *                      rf_chunk_pass3_x(x0, x1, ..., last = (s is the final range))
 * Same kernels on the same data as the plain chunked plan: the field is bit-identical. */
int rf_chunk_exchange_x(rf_ctx* ctx, int chunk /* -1: every chunk */, int x0, int x1, int wait_pass1);
int rf_chunk_pass2_x(rf_ctx* ctx, int chunk /* -1: every chunk */, int x0, int x1,
                     int wait_index /* of rf_chunk_arrived, -1: none */);
int rf_chunk_pass3_x(rf_ctx* ctx, int x0, int x1, const double* zscale, int want_stats, int last);

/* raw device pointers for zero-copy wrapping (torch / DLPack) and for the all-to-all */
void* rf_device_ptr(rf_ctx* ctx, int which);
long long rf_buffer_bytes(rf_ctx* ctx, int which);
/* 1 if the (nx, ny, nz, dtype) grid runs on the tile kernels, 0 if on the generic path */
int rf_is_fast(rf_ctx* ctx);
/* 1 if the hot path runs on the persistent TMA-pipelined kernels (float32, axes of 64..4096 points
 * and -- once rf_set_power has been called -- a sigma table the single-load evaluation covers),
 * 0 if pass 1 falls back to the rf_fast.cuh tile kernels or the grid is on the generic path */
int rf_is_pipelined(rf_ctx* ctx);
/* 0: passes 2 and 3 are separate kernels; 2: fused in k23q_fused_yz (whole-field rf_finish_yz only;
 * the chunked exchange runs them separately) */
int rf_is_fused(rf_ctx* ctx);
/* "k1;k2;k3": names of the kernels pass 1 / 2 / 3 launch for this context as configured now */
int rf_kernel_names(rf_ctx* ctx, char* buf, int buflen);
/* 1: a k-space block is pending (between rf_generate_xpass and rf_finish_yz) or nothing has been
 * generated; 2: RF_BUF_FIELD holds a real-space field */
int rf_field_state(rf_ctx* ctx);
/* number of kernels launched by this context so far */
long long rf_launch_count(rf_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* RF_B200_H */
