// rf_api.cu -- context management and the C ABI of librf_b200.so (see include/rf_b200.h).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../../include/rf_b200.h"
#include "rf_generic.cuh"
#include "rf_launch.h"
#include "rf_lens.cuh"

using namespace rf;

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define RF_CUDA(expr)                                                                     \
  do {                                                                                    \
    cudaError_t e__ = (cudaError_t)(expr);                                                \
    if (e__ != cudaSuccess)                                                               \
      return fail(RF_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));      \
  } while (0)

extern "C" const char* rf_last_error(void) { return g_err.c_str(); }
extern "C" int rf_version(void) { return 100; }

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
struct rf_ctx {
  int nx = 0, ny = 0, nz = 0, dtype = 0, rank = 0, world = 1;
  int nxl = 0, nyl = 0;     // local slab extents
  bool fast = false;
  bool pipe = false;        // TMA-pipelined kernels (f32, env RF_PIPE != 0)
  bool fuse = false;        // passes 2 + 3 in one kernel, k23q_fused_yz (env RF_FUSE: 0 off, 2 wherever it fits)
  bool k1q = true;          // pass 1 with two CTAs per SM (env RF_K1Q=0: one-CTA k1p)
  bool k1r = true;          // pass 1 with two ping-pong groups in one CTA per SM (env RF_K1R=0: k1q)
  bool k3q = true;          // pass 3 with the three-deep in-place ring (env RF_K3Q=0: k3p)
  bool pass2_done = false;  // pass 2 already ran per chunk (rf_chunk_pass2): rf_finish_yz starts at pass 3
  bool big = true;          // 2048 / 4096-point strided axes on the rf_big.cuh kernels (env RF_BIG=0: rf_fast.cuh)
  bool lens4 = true;        // register-resident lensing scan where nz allows (env RF_LENS4=0: generic scan)
  int env_lag = 0;          // env RF_LAG: planes between a plane's y pass and its z pass in k23q (0: automatic)
  int hinge_lo = 0, hinge_hi = -1;   // cells of the sigma table the grid's modes can touch
  unsigned int* sched = nullptr;   // [1 + nxl] ticket + per-plane counters of k23q
  void* peer_recv[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  bool direct = false;      // pass 1 stores straight into the peers' receive buffers
  // chunked exchange (rf_chunk_*): the packed z columns are cut into ck_n chunks; pass 1 of chunk
  // c writes send[c] = [nx][nyl][wc], copy engines push its G blocks into the peers' recv[c] =
  // [G][nxl][nyl][wc] while pass 1 of chunk c+1 and pass 2 of chunk c-1 run on the SMs
  int ck_n = 0;             // chunks (0: off)
  int ck_cur = -1;          // chunk the next pass-1 / pass-2 launch works on (-1: whole field)
  int ck_reserve = 0;       // SMs left out of the chunk launches' grids (world > 1; env RF_RESERVE_SMS)
  void* send = nullptr;     // [ck_n][nx][nyl][wc]
  cudaStream_t comm = nullptr;
  cudaEvent_t ev_k1 = nullptr, ev_x = nullptr, ev_s = nullptr;
  std::vector<cudaEvent_t> ev_xc;   // per chunk: exchange complete (recorded on the comm stream)
  // timing of a chunked step: one (start, stop, pass) record per launch, reused every step
  std::vector<cudaEvent_t> seg_a, seg_b;
  std::vector<int> seg_pass;
  int seg_used = 0;
  long long part_used = 0;          // statistics partials written by this step's pass-3 launches
  bool zs_loaded = false;           // this step's z-scale table is on the device
  bool last_chunked = false;        // the last step ran through rf_chunk_*
  int num_sms = 148;
  size_t esize = 4;         // sizeof(T)
  cudaStream_t own_stream = nullptr, stream = nullptr;
  void* field = nullptr;    // RF_BUF_FIELD
  void* recv = nullptr;     // RF_BUF_RECV (world > 1) or generic-path scratch
  size_t buf_bytes = 0;
  double* partials = nullptr;
  long long nparts = 0, parts_cap = 0;
  double* stats_dev = nullptr;
  bool stats_valid = false;
  // twiddles exp(+2 pi i k / n) per axis (+ nz/2 for the c2r half-length transform)
  void *tw_x = nullptr, *tw_y = nullptr, *tw_z = nullptr, *tw_zh = nullptr;
  // sigma table
  void* knots = nullptr;
  void* cells = nullptr;
  void* hinge = nullptr;     // float4 per cell (f32 only)
  void* plane = nullptr;     // [nx][nyl] packed iz = 0 / nz/2 column for k1p
  bool hinge_ok = false;     // every non-zero mode of the grid falls in a hinge-valid cell
  int ntab = 0, lut_n = 0, lut_base = 0, lut_shift = 0;
  double spacing = 0.0;
  bool have_power = false;
  double* noise = nullptr;
  void *zs_dev = nullptr, *da_dev = nullptr, *db_dev = nullptr;   // nz-length tables of T
  double* lens_dev = nullptr;   // 5 * nz coefficient tables of rf_lensing
  long long launches = 0;
  bool field_is_real = false;
  bool aux_is_psi = false;          // RF_BUF_AUX (= the receive buffer) holds the lensing potential
  cudaStream_t copy_stream = nullptr;   // streamed output (rf_copy_rows / rf_finish_yz_streamed)
  cudaEvent_t ev_copy = nullptr;
  std::vector<cudaEvent_t> ev_rng;      // per x range of a streamed finish: pass 3 of the range is done
  bool timing = false;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

static void mark(rf_ctx* c, int i) {
  if (c->timing) cudaEventRecord(c->ev[i], c->stream);
}

static bool is_pow2(int n) { return n > 0 && (n & (n - 1)) == 0; }
static bool fast_size(int n) { return is_pow2(n) && n >= 64 && n <= 4096; }

template <typename T>
static int upload_twiddles(void** dst, int n) {
  std::vector<cpx<T>> h(n);
  for (int k = 0; k < n; ++k) {
    // exact quadrant reduction through the argument of cos/sin(pi * x)
    long double a = 2.0L * (long double)k / (long double)n;
    h[k] = {(T)cosl(M_PIl * a), (T)sinl(M_PIl * a)};
  }
  // pin the exactly-representable points
  h[0] = {(T)1, (T)0};
  if (n % 2 == 0) h[n / 2] = {(T)-1, (T)0};
  if (n % 4 == 0) { h[n / 4] = {(T)0, (T)1}; h[3 * n / 4] = {(T)0, (T)-1}; }
  RF_CUDA(cudaMalloc(dst, sizeof(cpx<T>) * n));
  RF_CUDA(cudaMemcpy(*dst, h.data(), sizeof(cpx<T>) * n, cudaMemcpyHostToDevice));
  return 0;
}

extern "C" int rf_create(int nx, int ny, int nz, int dtype, int rank, int world, rf_ctx** out) {
  if (!out) return fail(RF_ERR_ARG, "rf_create: out is NULL");
  if (nx <= 0 || ny <= 0 || nz <= 0 || (nx % 2) || (ny % 2) || (nz % 2))
    return fail(RF_ERR_ARG, "All shape dimensions must be even.");
  if (dtype != RF_F32 && dtype != RF_F64) return fail(RF_ERR_ARG, "rf_create: bad dtype");
  if (world < 1 || rank < 0 || rank >= world) return fail(RF_ERR_ARG, "rf_create: bad rank/world");
  if (world > 1 && ((nx % world) || (ny % world)))
    return fail(RF_ERR_ARG, "rf_create: nx and ny must be divisible by world");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(RF_ERR_CUDA, "rf_create: no CUDA device available (this library has no CPU path)");
  rf_ctx* c = new rf_ctx();
  c->nx = nx; c->ny = ny; c->nz = nz; c->dtype = dtype; c->rank = rank; c->world = world;
  c->nxl = nx / world; c->nyl = ny / world;
  c->esize = dtype == RF_F32 ? 4 : 8;
  c->fast = fast_size(nx) && fast_size(ny) && fast_size(nz);
  if (world > 1 && !c->fast) {
    delete c;
    return fail(RF_ERR_ARG, "rf_create: slab decomposition needs power-of-two dims in [64, 4096]");
  }
  {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, dev);
    // run-time switches are read once, here
    auto env_off = [](const char* name) { const char* e = getenv(name); return e && e[0] == '0'; };
    c->pipe = c->fast && dtype == RF_F32 && !env_off("RF_PIPE");
    // Passes 2 and 3 in one kernel, the z pass reading the y-pass output from L2:
    // default where it was measured faster than k2p + k3q -- both axes <= 1024 and a field of 2^26
    // cells or more; RF_FUSE=2 (or 1): wherever it fits; RF_FUSE=0: separate kernels.
    const bool fuse_ok = c->pipe && ny <= 1024 && nz <= 2048 && nz >= 128;
    const long long cells_local = (long long)(nx / world) * ny * nz;
    if (const char* envf = getenv("RF_FUSE")) c->fuse = fuse_ok && (envf[0] == '1' || envf[0] == '2');
    else c->fuse = fuse_ok && nz <= 1024 && cells_local >= (1LL << 26);
    c->k1q = !env_off("RF_K1Q");
    c->k1r = !env_off("RF_K1R");
    c->k3q = !env_off("RF_K3Q");
    c->big = !env_off("RF_BIG");
    c->lens4 = !env_off("RF_LENS4");
    if (const char* el = getenv("RF_LAG")) c->env_lag = std::max(0, atoi(el));
  }
  RF_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  c->stream = c->own_stream;
  // local k-space block in the public layout is the largest thing we ever hold
  const size_t nzh = nz / 2 + 1;
  c->buf_bytes = (size_t)nx * c->nyl * nzh * 2 * c->esize;
  RF_CUDA(cudaMalloc(&c->field, c->buf_bytes));
  if (world > 1) RF_CUDA(cudaMalloc(&c->recv, c->buf_bytes));
  RF_CUDA(cudaMalloc(&c->stats_dev, 2 * sizeof(double)));
  int e = 0;
  if (dtype == RF_F32) {
    e |= upload_twiddles<float>(&c->tw_x, nx);
    e |= upload_twiddles<float>(&c->tw_y, ny);
    e |= upload_twiddles<float>(&c->tw_z, nz);
    e |= upload_twiddles<float>(&c->tw_zh, nz / 2);
  } else {
    e |= upload_twiddles<double>(&c->tw_x, nx);
    e |= upload_twiddles<double>(&c->tw_y, ny);
    e |= upload_twiddles<double>(&c->tw_z, nz);
    e |= upload_twiddles<double>(&c->tw_zh, nz / 2);
  }
  if (e) { return RF_ERR_CUDA; }
  RF_CUDA(cudaMalloc(&c->zs_dev, nz * c->esize));
  RF_CUDA(cudaMalloc(&c->da_dev, nz * c->esize));
  RF_CUDA(cudaMalloc(&c->db_dev, nz * c->esize));
  *out = c;
  return RF_OK;
}

extern "C" int rf_destroy(rf_ctx* c) {
  if (!c) return RF_OK;
  cudaStreamSynchronize(c->stream);
  void* ptrs[] = {c->field, c->recv, c->partials, c->stats_dev, c->tw_x, c->tw_y, c->tw_z,
                  c->tw_zh, c->knots, c->cells, c->hinge, c->plane, c->sched, c->lens_dev, c->noise, c->zs_dev, c->da_dev, c->db_dev};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  if (c->send) cudaFree(c->send);
  for (cudaEvent_t ev : c->seg_a) cudaEventDestroy(ev);
  for (cudaEvent_t ev : c->seg_b) cudaEventDestroy(ev);
  for (cudaEvent_t ev : c->ev_xc) cudaEventDestroy(ev);
  if (c->comm) cudaStreamDestroy(c->comm);
  for (cudaEvent_t ev : c->ev_rng) cudaEventDestroy(ev);
  if (c->ev_copy) cudaEventDestroy(c->ev_copy);
  if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
  for (cudaEvent_t e : {c->ev_k1, c->ev_x, c->ev_s})
    if (e) cudaEventDestroy(e);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  delete c;
  return RF_OK;
}

extern "C" int rf_set_stream(rf_ctx* c, void* s) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  c->stream = s ? (cudaStream_t)s : c->own_stream;
  return RF_OK;
}
extern "C" int rf_sync(rf_ctx* c) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  RF_CUDA(cudaStreamSynchronize(c->stream));
  return RF_OK;
}
extern "C" void* rf_host_alloc(long long bytes) {
  void* p = nullptr;
  if (bytes <= 0) return nullptr;
  if (cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocDefault) != cudaSuccess) {
    g_err = "rf_host_alloc: cudaHostAlloc failed";
    return nullptr;
  }
  return p;
}
extern "C" void rf_host_free(void* p) {
  if (p) cudaFreeHost(p);
}
extern "C" int rf_set_timing(rf_ctx* c, int enabled) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  if (enabled && !c->ev[0])
    for (int i = 0; i < 6; ++i) RF_CUDA(cudaEventCreate(&c->ev[i]));
  c->timing = enabled != 0;
  return RF_OK;
}
extern "C" int rf_get_timing(rf_ctx* c, double* a, double* b, double* d) {
  if (!c || !c->timing || !c->fast) return fail(RF_ERR_STATE, "rf_get_timing: timing is off or generic path");
  RF_CUDA(cudaStreamSynchronize(c->stream));
  float t[3] = {0, 0, 0};
  if (c->last_chunked && c->seg_used > 0) {
    // chunked step: the passes are interleaved; report the sum of each pass's launches
    for (int k = 0; k < c->seg_used; ++k) {
      float dt = 0;
      RF_CUDA(cudaEventElapsedTime(&dt, c->seg_a[k], c->seg_b[k]));
      t[c->seg_pass[k]] += dt;
    }
  } else {
    for (int i = 0; i < 3; ++i) RF_CUDA(cudaEventElapsedTime(&t[i], c->ev[2 * i], c->ev[2 * i + 1]));
  }
  if (a) *a = t[0];
  if (b) *b = t[1];
  if (d) *d = t[2];
  return RF_OK;
}
// ------------------------------------------------------------------------------------------
// direct exchange over peer memory
// ------------------------------------------------------------------------------------------
extern "C" int rf_ipc_export(rf_ctx* c, int which, unsigned char* handle64) {
  if (!c || !handle64) return fail(RF_ERR_ARG, "rf_ipc_export: null argument");
  void* p = rf_device_ptr(c, which);
  if (!p) return fail(RF_ERR_STATE, "rf_ipc_export: buffer not allocated");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  cudaIpcMemHandle_t h;
  RF_CUDA(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, 64);
  return RF_OK;
}
extern "C" int rf_ipc_open(const unsigned char* handle64, void** ptr) {
  if (!handle64 || !ptr) return fail(RF_ERR_ARG, "rf_ipc_open: null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  RF_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return RF_OK;
}
extern "C" int rf_ipc_close(void* ptr) {
  if (ptr) RF_CUDA(cudaIpcCloseMemHandle(ptr));
  return RF_OK;
}
extern "C" int rf_set_peer_recv(rf_ctx* c, int rank, void* recv_ptr) {
  if (!c || rank < 0 || rank >= c->world || c->world > 8)
    return fail(RF_ERR_ARG, "rf_set_peer_recv: bad rank (direct exchange supports up to 8 ranks)");
  c->peer_recv[rank] = recv_ptr;
  return RF_OK;
}
extern "C" int rf_set_direct(rf_ctx* c, int enabled) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  if (enabled) {
    if (c->world < 2 || c->world > 8 || !is_pow2(c->nxl))
      return fail(RF_ERR_ARG, "rf_set_direct: needs 2..8 ranks and a power-of-two x-slab");
    c->peer_recv[c->rank] = c->recv;
    for (int r = 0; r < c->world; ++r)
      if (!c->peer_recv[r]) return fail(RF_ERR_STATE, "rf_set_direct: a peer receive buffer is not set");
  }
  c->direct = enabled != 0;
  return RF_OK;
}

extern "C" int rf_is_fast(rf_ctx* c) { return c && c->fast ? 1 : 0; }
// pass 1 runs on the TMA-pipelined generator only when the sigma table allows the single-load
// hinge evaluation (build_table: hinge_ok) and, for 2048 / 4096-point x axes, when the cell table
// fits in shared memory; otherwise the rf_fast.cuh kernel generates and the x pass
static bool k1_is_pipelined(rf_ctx* c) {
  bool k1_pipe = c->pipe && c->nx <= 4096 && c->hinge_ok;
  if (k1_pipe && c->nx > 1024) {
    k1_pipe = c->big && (c->nx == 2048 ? PipeLaunch<2048>::k1p_fits(c->noise ? 0 : c->lut_n)
                                        : PipeLaunch<4096>::k1p_fits(c->noise ? 0 : c->lut_n));
  }
  return k1_pipe;
}
extern "C" int rf_is_pipelined(rf_ctx* c) {
  if (!c || !c->pipe || c->nx > 4096 || c->ny > 4096 || c->nz > 4096) return 0;
  return (!c->have_power || k1_is_pipelined(c)) ? 1 : 0;
}
extern "C" int rf_field_state(rf_ctx* c) { return !c ? 0 : (c->field_is_real ? 2 : 1); }
static const char* k1_small_name(rf_ctx* c) {
  bool q = false, r = false;
  switch (c->nx) {
    case 64: q = PipeLaunch<64>::k1q_fits(c->lut_n); r = PipeLaunch<64>::k1r_fits(c->lut_n); break;
    case 128: q = PipeLaunch<128>::k1q_fits(c->lut_n); r = PipeLaunch<128>::k1r_fits(c->lut_n); break;
    case 256: q = PipeLaunch<256>::k1q_fits(c->lut_n); r = PipeLaunch<256>::k1r_fits(c->lut_n); break;
    case 512: q = PipeLaunch<512>::k1q_fits(c->lut_n); r = PipeLaunch<512>::k1r_fits(c->lut_n); break;
    case 1024: q = PipeLaunch<1024>::k1q_fits(c->lut_n); r = PipeLaunch<1024>::k1r_fits(c->lut_n); break;
  }
  if (c->k1r && r) return "k1r_generate_xfft";
  return (c->k1q && q) ? "k1q_generate_xfft" : "k1p_generate_xfft";
}
// "k1;k2;k3" -- the kernels rf_generate_xpass / rf_finish_yz launch for this context as configured
// now (after rf_set_power); a fused y + z pass reports the same name for k2 and k3
extern "C" int rf_kernel_names(rf_ctx* c, char* buf, int buflen) {
  if (!c || !buf || buflen < 8) return fail(RF_ERR_ARG, "rf_kernel_names: bad argument");
  std::string k1, k2, k3;
  if (!c->fast) {
    k1 = "gen_kspace_public_kernel+generic_line_fft_kernel"; k2 = k3 = "generic_line_fft_kernel";
  } else {
    const bool f32 = c->dtype == RF_F32;
    if (f32 && c->have_power && k1_is_pipelined(c))
      k1 = c->nx > 1024 ? "k1b_generate_xfft" : k1_small_name(c);
    else k1 = "k1_generate_xfft";
    if (f32 && c->pipe && (c->ny <= 1024 || (c->ny <= 4096 && c->big)))
      k2 = c->ny > 1024 ? "k2b_strided_fft" : "k2p_strided_fft";
    else k2 = "k2_strided_fft";
    if (f32 && c->pipe && c->nz <= 4096) k3 = (c->k3q && c->nz >= 128) ? "k3q_c2r_lines" : "k3p_c2r_lines";   // (nz = 64: the ring does not fit)
    else k3 = "k3_c2r_lines";
    if (c->fuse && k2 == "k2p_strided_fft" && (long long)c->ny * (c->nz / 2) * 8 >= pipe::kTileBytes)
      k2 = k3 = "k23q_fused_yz";
  }
  const std::string all = k1 + ";" + k2 + ";" + k3;
  if ((int)all.size() + 1 > buflen) return fail(RF_ERR_ARG, "rf_kernel_names: buffer too small");
  memcpy(buf, all.c_str(), all.size() + 1);
  return RF_OK;
}
extern "C" int rf_is_fused(rf_ctx* c) { return c && c->fuse ? 2 : 0; }
extern "C" long long rf_launch_count(rf_ctx* c) { return c ? c->launches : 0; }
extern "C" long long rf_field_elems(rf_ctx* c) {
  return c ? (long long)c->nxl * c->ny * c->nz : 0;
}
extern "C" void* rf_device_ptr(rf_ctx* c, int which) {
  if (!c) return nullptr;
  switch (which) {
    case RF_BUF_FIELD: return c->field;
    case RF_BUF_RECV: return c->recv;
    case RF_BUF_PARTIALS: return c->partials;
  }
  return nullptr;
}
extern "C" long long rf_buffer_bytes(rf_ctx* c, int which) {
  if (!c) return 0;
  switch (which) {
    case RF_BUF_FIELD: return (long long)c->buf_bytes;
    case RF_BUF_RECV: return c->recv ? (long long)c->buf_bytes : 0;
    case RF_BUF_PARTIALS: return c->nparts * 2 * (long long)sizeof(double);
  }
  return 0;
}

// ------------------------------------------------------------------------------------------
// sigma table
// ------------------------------------------------------------------------------------------
template <typename T> struct FloatBits;
template <> struct FloatBits<float> { static constexpr int kMant = 23; };   // bits below the exponent
template <> struct FloatBits<double> { static constexpr int kMant = 20; };  // ... in the high word

static int cell_of(float v, int shift) {
  int32_t b;
  memcpy(&b, &v, 4);
  return b >> shift;
}
static int cell_of(double v, int shift) {
  int64_t b;
  memcpy(&b, &v, 8);
  return (int)((int32_t)(b >> 32) >> shift);
}

// Builds the knot array and the direct-indexed cell table described in rf_common.cuh.
template <typename T>
static int build_table(rf_ctx* c, int n, const double* k, const double* sigma) {
  const double N3 = (double)c->nx * c->ny * c->nz;
  std::vector<knot<T>> kn(n);
  std::vector<double> l2(n), sg(n), slope(n, 0.0);
  for (int j = 0; j < n; ++j) {
    kn[j].ksq = (T)(k[j] * k[j]);
    l2[j] = 2.0 * std::log2(k[j]);
    sg[j] = sigma[j] / N3;
  }
  for (int j = 0; j < n; ++j) {
    if (j + 1 < n) slope[j] = (sg[j + 1] - sg[j]) / (l2[j + 1] - l2[j]);
    kn[j].l2 = (T)l2[j];
    kn[j].sig = (T)sg[j];
    kn[j].slope = (T)slope[j];
    if (j > 0 && !(kn[j].ksq > kn[j - 1].ksq))
      return fail(RF_ERR_ARG, "Power spectrum k values are not strictly increasing.");
  }
  // mantissa bits per cell: the fewest (4..8) that leave at most one knot per cell
  int mb = 4, shift = 0, base = 0, ln = 0;
  for (;; ++mb) {
    shift = FloatBits<T>::kMant - mb;
    base = cell_of(kn[0].ksq, shift) - 1;
    ln = cell_of(kn[n - 1].ksq, shift) - base + 2;
    int worst = 1;
    for (int j = 1, run = 1; j < n; ++j) {
      run = (cell_of(kn[j].ksq, shift) == cell_of(kn[j - 1].ksq, shift)) ? run + 1 : 1;
      worst = std::max(worst, run);
    }
    if (worst <= 1 || mb == 8) break;
  }
  if (ln < 3 || ln > (1 << 22)) return fail(RF_ERR_ARG, "rf_set_power: bad k range");
  const T inf = std::numeric_limits<T>::infinity();
  auto zero_cell = [&](int bin) {
    sigcell<T> z;
    z.c0a = z.c1a = z.c0b = z.c1b = (T)0; z.t = inf; z.bin_a = (T)bin; z.slow = (T)0; z.pad = (T)0;
    return z;
  };
  // sigma = c0 + c1*lg on segment j, anchored at the segment midpoint so that the rounding of
  // c1 to T costs nothing there
  auto seg_coef = [&](int j, T& c0, T& c1) {
    c1 = (T)slope[j];
    const double mid = 0.5 * (l2[j] + l2[j + 1]);
    c0 = (T)((sg[j] + slope[j] * (mid - l2[j])) - (double)c1 * mid);
  };
  std::vector<sigcell<T>> cells(ln);
  cells[0] = zero_cell(-1);
  cells[ln - 1] = zero_cell(n - 1);
  int next = 0;   // first knot not below the current cell
  for (int e = 1; e < ln - 1; ++e) {
    const int id = base + e;
    while (next < n && cell_of(kn[next].ksq, shift) < id) ++next;
    int cnt = 0;
    while (next + cnt < n && cell_of(kn[next + cnt].ksq, shift) == id) ++cnt;
    const int jlo = next - 1;
    sigcell<T> s = zero_cell(jlo);
    if (cnt == 0) {
      seg_coef(jlo, s.c0a, s.c1a);
      s.c0b = s.c0a; s.c1b = s.c1a;
    } else if (cnt == 1 && next < n - 1) {
      if (jlo >= 0) seg_coef(jlo, s.c0a, s.c1a);
      seg_coef(next, s.c0b, s.c1b);
      s.t = kn[next].ksq;
    } else {
      s.slow = (T)1;
    }
    cells[e] = s;
  }
  // hinge form of the same cells (float path): segment below the knot + slope change at it
  std::vector<float> hinge;
  bool hinge_ok = false;
  if (std::is_same<T, float>::value) {
    hinge.assign((size_t)4 * ln, 0.f);
    std::vector<char> valid(ln, 0);
    next = 0;
    for (int e = 1; e < ln - 1; ++e) {
      const int id = base + e;
      while (next < n && cell_of(kn[next].ksq, shift) < id) ++next;
      int cnt = 0;
      while (next + cnt < n && cell_of(kn[next + cnt].ksq, shift) == id) ++cnt;
      const int jlo = next - 1;
      if (jlo < 0 || cnt > 1 || next + cnt >= n) continue;   // touches the table's ends, or 2+ knots
      float c0, c1;
      {
        T a, b;
        seg_coef(jlo, a, b);
        c0 = (float)a; c1 = (float)b;
      }
      float lt = 0.f, d = 0.f;
      if (cnt == 1) { lt = (float)l2[next]; d = (float)(slope[next] - slope[jlo]); }
      hinge[4 * e + 0] = c0; hinge[4 * e + 1] = c1; hinge[4 * e + 2] = lt; hinge[4 * e + 3] = d;
      valid[e] = 1;
    }
    // grid modes: smallest non-zero k^2 and largest k^2 (with a cell of margin either side)
    const double two_pi = 6.283185307179586476925286766559;
    double kmin2 = 1e300, kmax2 = 0;
    for (int n_ax : {c->nx, c->ny, c->nz}) {
      double k0 = two_pi / ((double)n_ax * c->spacing);
      kmin2 = std::min(kmin2, k0 * k0);
      kmax2 += (0.5 * n_ax * k0) * (0.5 * n_ax * k0);
    }
    const int elo = cell_of((T)kmin2, shift) - base - 1, ehi = cell_of((T)kmax2, shift) - base + 1;
    hinge_ok = elo >= 1 && ehi <= ln - 2;
    for (int e = std::max(elo, 1); hinge_ok && e <= ehi; ++e) hinge_ok = valid[e] != 0;
    c->hinge_lo = std::max(elo, 0);
    c->hinge_hi = std::min(ehi, ln - 1);
  }
  if (c->knots) cudaFree(c->knots);
  if (c->cells) cudaFree(c->cells);
  if (c->hinge) cudaFree(c->hinge);
  c->knots = nullptr; c->cells = nullptr; c->hinge = nullptr;
  c->hinge_ok = false;
  if (!hinge.empty()) {
    RF_CUDA(cudaMalloc(&c->hinge, sizeof(float) * hinge.size()));
    RF_CUDA(cudaMemcpy(c->hinge, hinge.data(), sizeof(float) * hinge.size(), cudaMemcpyHostToDevice));
    c->hinge_ok = hinge_ok;
  }
  RF_CUDA(cudaMalloc(&c->knots, sizeof(knot<T>) * n));
  RF_CUDA(cudaMalloc(&c->cells, sizeof(sigcell<T>) * ln));
  // the table may be replaced while earlier work is still queued on the stream
  RF_CUDA(cudaStreamSynchronize(c->stream));
  RF_CUDA(cudaMemcpy(c->knots, kn.data(), sizeof(knot<T>) * n, cudaMemcpyHostToDevice));
  RF_CUDA(cudaMemcpy(c->cells, cells.data(), sizeof(sigcell<T>) * ln, cudaMemcpyHostToDevice));
  c->ntab = n; c->lut_n = ln; c->lut_base = base; c->lut_shift = shift;
  return RF_OK;
}

extern "C" int rf_set_power(rf_ctx* c, int n, const double* k, const double* sigma, double spacing) {
  if (!c || !k || !sigma) return fail(RF_ERR_ARG, "rf_set_power: null argument");
  if (n < 2 || n > 32000) return fail(RF_ERR_ARG, "rf_set_power: need 2 <= n <= 32000 knots");
  if (!(spacing > 0)) return fail(RF_ERR_ARG, "rf_set_power: spacing must be > 0");
  for (int j = 0; j < n; ++j) {
    if (!std::isfinite(k[j]) || !(k[j] > 0)) return fail(RF_ERR_ARG, "Power spectrum min(k) is <= 0.");
    if (!std::isfinite(sigma[j]) || sigma[j] < 0)
      return fail(RF_ERR_ARG, "Power values P(k) are not all non-negative.");
  }
  c->spacing = spacing;
  int e = c->dtype == RF_F32 ? build_table<float>(c, n, k, sigma) : build_table<double>(c, n, k, sigma);
  c->have_power = (e == RF_OK);
  return e;
}

// rows x in [x0, x1) of the white-noise array [nx][ny][nz/2+1][2] (contiguous on host and device):
// the caller may draw and upload the deviates slab by slab without ever holding 16 B per mode on
// the host (random.py:25-27 notes that the reference's temporary "should" be chunked)
extern "C" int rf_set_noise_rows(rf_ctx* c, const double* host_rows, int x0, int x1) {
  if (!c || !host_rows) return fail(RF_ERR_ARG, "rf_set_noise_rows: null argument");
  if (c->world > 1) return fail(RF_ERR_ARG, "rf_set_noise: single-GPU only");
  if (x0 < 0 || x1 > c->nx || x0 >= x1) return fail(RF_ERR_ARG, "rf_set_noise_rows: bad x range");
  const size_t row = (size_t)2 * c->ny * (c->nz / 2 + 1);
  if (!c->noise) {
    RF_CUDA(cudaStreamSynchronize(c->stream));
    RF_CUDA(cudaMalloc((void**)&c->noise, row * c->nx * sizeof(double)));
  }
  // ordered after the work already queued on the stream (a previous step may still read the array);
  // the source is pageable in general, so the call returns once the rows are staged
  RF_CUDA(cudaMemcpyAsync(c->noise + row * x0, host_rows, row * (x1 - x0) * sizeof(double), cudaMemcpyHostToDevice,
                          c->stream));
  RF_CUDA(cudaStreamSynchronize(c->stream));
  return RF_OK;
}

extern "C" int rf_set_noise(rf_ctx* c, const double* host_noise) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  if (!host_noise) {
    RF_CUDA(cudaStreamSynchronize(c->stream));
    if (c->noise) cudaFree(c->noise);
    c->noise = nullptr;
    return RF_OK;
  }
  // uploaded in slabs of about 256 MB
  const size_t row = (size_t)2 * c->ny * (c->nz / 2 + 1);
  const int step = (int)std::max<size_t>(1, ((size_t)32 << 20) / row);
  for (int x0 = 0; x0 < c->nx; x0 += step) {
    int e = rf_set_noise_rows(c, host_noise + row * x0, x0, std::min(c->nx, x0 + step));
    if (e) return e;
  }
  return RF_OK;
}

// ------------------------------------------------------------------------------------------
// TMA tensor maps (driver entry point fetched through the runtime; no -lcuda link)
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &p, 12000, cudaEnableDefault, &q) ==
            cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}
// float32 tensor of `rank` dims (fastest first); strides[d-1] = byte stride of dim d
static int make_tensor_map(CUtensorMap* m, void* base, int rank, const uint64_t* dims,
                           const uint64_t* strides, const uint32_t* box) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(RF_ERR_CUDA, "cuTensorMapEncodeTiled is not available");
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int d = 0; d < rank; ++d) { gd[d] = dims[d]; bx[d] = box[d]; es[d] = 1; }
  for (int d = 0; d + 1 < rank; ++d) gs[d] = strides[d];
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, base, gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(RF_ERR_CUDA, "cuTensorMapEncodeTiled failed: " + std::to_string((int)r));
  return RF_OK;
}

template <typename T>
static ModeGeom<T> make_geom(rf_ctx* c, uint64_t seed, int potential) {
  ModeGeom<T> g;
  g.nx = c->nx; g.ny = c->ny; g.nz = c->nz;
  const double two_pi = 6.283185307179586476925286766559;
  auto k0sq = [&](int n) {
    double k0 = two_pi / ((double)n * c->spacing);
    return (T)(k0 * k0);
  };
  g.k0x2 = k0sq(c->nx); g.k0y2 = k0sq(c->ny); g.k0z2 = k0sq(c->nz);
  g.key0 = (uint32_t)seed; g.key1 = (uint32_t)(seed >> 32);
  g.noise = c->noise;
  g.potential = potential;
  return g;
}
template <typename T>
static SigmaTable<T> make_tab(rf_ctx* c) {
  SigmaTable<T> t;
  t.knots = (const knot<T>*)c->knots; t.cells = (const sigcell<T>*)c->cells; t.hinge = c->hinge;
  t.n = c->ntab; t.lut_n = c->lut_n; t.lut_base = c->lut_base; t.shift = c->lut_shift;
  return t;
}

static int ensure_partials(rf_ctx* c, long long n) {
  if (n > c->parts_cap) {
    if (c->partials) cudaFree(c->partials);
    c->partials = nullptr;
    RF_CUDA(cudaMalloc((void**)&c->partials, sizeof(double) * 2 * n));
    c->parts_cap = n;
  }
  c->nparts = n;
  return RF_OK;
}

// ------------------------------------------------------------------------------------------
// dispatch over the compiled tile-kernel sizes
// ------------------------------------------------------------------------------------------
#define RF_DISPATCH_PIPE(N_, ...)                       \
  switch (N_) {                                         \
    case 64: { constexpr int NN = 64; __VA_ARGS__; } break;    \
    case 128: { constexpr int NN = 128; __VA_ARGS__; } break;  \
    case 256: { constexpr int NN = 256; __VA_ARGS__; } break;  \
    case 512: { constexpr int NN = 512; __VA_ARGS__; } break;  \
    case 1024: { constexpr int NN = 1024; __VA_ARGS__; } break;\
    case 2048: { constexpr int NN = 2048; __VA_ARGS__; } break;\
    case 4096: { constexpr int NN = 4096; __VA_ARGS__; } break;\
    default: return fail(RF_ERR_ARG, "size not compiled");\
  }

#define RF_DISPATCH_PIPE1K(N_, ...)                     \
  switch (N_) {                                         \
    case 64: { constexpr int NN = 64; __VA_ARGS__; } break;    \
    case 128: { constexpr int NN = 128; __VA_ARGS__; } break;  \
    case 256: { constexpr int NN = 256; __VA_ARGS__; } break;  \
    case 512: { constexpr int NN = 512; __VA_ARGS__; } break;  \
    case 1024: { constexpr int NN = 1024; __VA_ARGS__; } break;\
    default: return fail(RF_ERR_ARG, "size not compiled");\
  }

#define RF_DISPATCH_N(N_, ...)                          \
  switch (N_) {                                         \
    case 64: { constexpr int NN = 64; __VA_ARGS__; } break;    \
    case 128: { constexpr int NN = 128; __VA_ARGS__; } break;  \
    case 256: { constexpr int NN = 256; __VA_ARGS__; } break;  \
    case 512: { constexpr int NN = 512; __VA_ARGS__; } break;  \
    case 1024: { constexpr int NN = 1024; __VA_ARGS__; } break;\
    case 2048: { constexpr int NN = 2048; __VA_ARGS__; } break;\
    case 4096: { constexpr int NN = 4096; __VA_ARGS__; } break;\
    default: return fail(RF_ERR_ARG, "size not compiled");\
  }

template <typename T>
static int run_generic_axis(rf_ctx* c, const void* in, void* out, int mode, int inverse, int n,
                            long long na, long long nb, long long sa_in, long long sb_in,
                            long long sa_out, long long sb_out, long long st_in, long long st_out,
                            T scale, cudaStream_t s) {
  LineSet ls;
  ls.na = na; ls.nb = nb; ls.sa_in = sa_in; ls.sb_in = sb_in; ls.sa_out = sa_out; ls.sb_out = sb_out;
  ls.stride_in = st_in; ls.stride_out = st_out; ls.n = n;
  size_t smem = (size_t)3 * n * sizeof(cpx<T>);
  if (smem > 220 * 1024) return fail(RF_ERR_ARG, "generic FFT: axis too long");
  if (smem > 48 * 1024)
    RF_CUDA(cudaFuncSetAttribute(generic_line_fft_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long lines = na * nb;
  int grid = (int)std::min<long long>(lines, 148LL * 8);
  int nt = n >= 512 ? 256 : (n >= 128 ? 128 : 64);
  generic_line_fft_kernel<T><<<grid, nt, smem, s>>>(in, out, ls, mode, inverse, scale);
  if (c) c->launches++;
  RF_CUDA(cudaGetLastError());
  return RF_OK;
}

// ------------------------------------------------------------------------------------------
// pass 1
// ------------------------------------------------------------------------------------------
template <typename T>
static int xpass_impl(rf_ctx* c, uint64_t seed, int potential) {
  ModeGeom<T> g = make_geom<T>(c, seed, potential);
  SigmaTable<T> tab = make_tab<T>(c);
  c->field_is_real = false;
  c->stats_valid = false;
  c->pass2_done = false;
  if constexpr (std::is_same<T, float>::value) {
    // (rf_big.cuh kernels stage the sigma cell table in shared memory; 32-byte-wide tiles
    // (nx = 4096) need nz/2 % 4 == 0, 64-byte ones nz/2 % 8 == 0 -- true for nz >= 64)
    const bool k1_pipe = k1_is_pipelined(c);
    if (k1_pipe) {
      pipe::K1PParams q;
      q.geom = g; q.tab = tab; q.tw = (const cpx<float>*)c->tw_x;
      q.ny_local = c->nyl; q.y0 = c->rank * c->nyl;
      const int nzc = c->nz / 2;
      if (!c->plane) RF_CUDA(cudaMalloc(&c->plane, sizeof(cpx<float>) * (size_t)c->nx * c->nyl));
      if (c->ck_cur <= 0) {
        const int total = c->nx * c->nyl;
        gen_packed_plane_kernel<float><<<std::min((total + 127) / 128, 148 * 16), 128, 0, c->stream>>>(
            g, tab, (cpx<float>*)c->plane, c->rank * c->nyl, c->nyl);
        c->launches++;
        RF_CUDA(cudaGetLastError());
      }
      q.plane = (const cpx<float>*)c->plane;
      q.z_base = 0;
      int e = 0;
      pipe::K1PMaps mout;
      memset(&mout, 0, sizeof(mout));
      c->pass2_done = false;
      if (c->ck_cur >= 0) {
        // one chunk of columns into its contiguous block of the send buffer
        const int wc = nzc / c->ck_n;
        q.z_base = c->ck_cur * wc;
        q.nxl = c->nxl;
        // one store map per destination block [nxl][nyl][wc]: blocks of the other ranks go to the
        // send buffer (the copy engines push them), this rank's own block goes straight into its
        // receive buffer -- no local copy
        const size_t block = (size_t)c->nxl * c->nyl * wc * sizeof(cpx<float>);
        const size_t chunk_off = (size_t)c->ck_cur * c->world * block;
        RF_DISPATCH_PIPE(c->nx, {
          const int tz = PipeLaunch<NN>::tile_cols();
          q.ztiles = wc / tz;
          q.ntiles = c->nyl * q.ztiles;
          uint64_t sout[2] = {(uint64_t)wc * 8, (uint64_t)c->nyl * wc * 8};
          uint64_t dout[3] = {(uint64_t)2 * wc, (uint64_t)c->nyl, (uint64_t)c->nxl};
          uint32_t bout[3] = {(uint32_t)2 * tz, 1, (uint32_t)std::min(PipeLaunch<NN>::box_rows(), c->nxl)};
          for (int d = 0; d < c->world; ++d) {
            char* base = (c->world > 1 && d == c->rank ? (char*)c->recv : (char*)c->send) + chunk_off + (size_t)d * block;
            int e2 = make_tensor_map(&mout.m[d], base, 3, dout, sout, bout);
            if (e2) return e2;
          }
          const bool pingpong = c->k1r && PipeLaunch<NN>::k1r_fits(c->lut_n);
          const bool twocta = !pingpong && c->k1q && PipeLaunch<NN>::k1q_fits(c->lut_n);
          // (ck_reserve: SMs left free for the 1-CTA fence collectives between chunks)
          int grid = std::min(q.ntiles, (twocta ? 2 : (pingpong ? 1 : PipeLaunch<NN>::ctas_per_sm())) * (c->num_sms - c->ck_reserve));
          e = pingpong ? PipeLaunch<NN>::k1r(c->noise != nullptr, potential != 0, mout, q, grid, c->stream)
              : twocta ? PipeLaunch<NN>::k1q(c->noise != nullptr, potential != 0, mout, q, grid, c->stream)
                       : PipeLaunch<NN>::k1p(c->noise != nullptr, potential != 0, mout, q, grid, c->stream);
        });
        c->launches++;
        RF_CUDA(e);
        return RF_OK;
      }
      RF_DISPATCH_PIPE(c->nx, {
        const int tz = PipeLaunch<NN>::tile_cols();
        q.ztiles = nzc / tz;
        q.ntiles = c->nyl * q.ztiles;
        uint64_t sout[2] = {(uint64_t)nzc * 8, (uint64_t)c->nyl * nzc * 8};
        if (c->direct) {
          // one map per destination: this rank's source block in that rank's receive buffer
          q.nxl = c->nxl;
          uint64_t dout[3] = {(uint64_t)2 * nzc, (uint64_t)c->nyl, (uint64_t)c->nxl};
          uint32_t bout[3] = {(uint32_t)2 * tz, 1, (uint32_t)std::min(PipeLaunch<NN>::box_rows(), c->nxl)};
          const size_t block = (size_t)c->nxl * c->nyl * nzc * sizeof(cpx<float>);
          for (int d = 0; d < c->world; ++d) {
            int e2 = make_tensor_map(&mout.m[d], (char*)c->peer_recv[d] + (size_t)c->rank * block, 3, dout, sout, bout);
            if (e2) return e2;
          }
        } else {
          q.nxl = c->nx;
          uint64_t dout[3] = {(uint64_t)2 * nzc, (uint64_t)c->nyl, (uint64_t)c->nx};
          uint32_t bout[3] = {(uint32_t)2 * tz, 1, (uint32_t)PipeLaunch<NN>::box_rows()};
          int e2 = make_tensor_map(&mout.m[0], c->field, 3, dout, sout, bout);
          if (e2) return e2;
        }
        const bool pingpong = c->k1r && PipeLaunch<NN>::k1r_fits(c->lut_n);
        const bool twocta = !pingpong && c->k1q && PipeLaunch<NN>::k1q_fits(c->lut_n);
        int grid = std::min(q.ntiles, (twocta ? 2 : (pingpong ? 1 : PipeLaunch<NN>::ctas_per_sm())) * c->num_sms);
        mark(c, 0);
        e = pingpong ? PipeLaunch<NN>::k1r(c->noise != nullptr, potential != 0, mout, q, grid, c->stream)
            : twocta ? PipeLaunch<NN>::k1q(c->noise != nullptr, potential != 0, mout, q, grid, c->stream)
                     : PipeLaunch<NN>::k1p(c->noise != nullptr, potential != 0, mout, q, grid, c->stream);
        mark(c, 1);
      });
      c->launches++;
      RF_CUDA(e);
      return RF_OK;
    }
  }
  if (c->fast) {
    K1Params<T> p;
    p.geom = g; p.tab = tab; p.out = (cpx<T>*)c->field; p.tw = (const cpx<T>*)c->tw_x;
    p.ny_local = c->nyl; p.y0 = c->rank * c->nyl;
    p.direct = c->direct ? 1 : 0;
    p.nxl_shift = 0;
    for (int d = 0; d < 8; ++d) p.peer[d] = nullptr;
    if (c->direct) {
      while ((1 << p.nxl_shift) < c->nxl) ++p.nxl_shift;
      const size_t block = (size_t)c->nxl * c->nyl * (c->nz / 2);
      for (int d = 0; d < c->world; ++d) p.peer[d] = (cpx<T>*)c->peer_recv[d] + (size_t)c->rank * block;
    }
    int e = 0;
    RF_DISPATCH_N(c->nx, {
      int grid = c->nyl * ((c->nz / 2) / FftPlan<T, NN>::TZ);
      mark(c, 0);
      e = Launch<T, NN>::k1(c->noise != nullptr, p, grid, c->stream);
      mark(c, 1);
    });
    c->launches++;
    RF_CUDA(e);
    return RF_OK;
  }
  // generic path: public-layout k-space, internal normalisation, then the x lines
  const int nzh = c->nz / 2 + 1;
  const long long total = (long long)c->nx * c->ny * nzh;
  int grid = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
  gen_kspace_public_kernel<T><<<grid, 256, 0, c->stream>>>(g, tab, (cpx<T>*)c->field, nullptr, (T)1, 0, c->ny);
  c->launches++;
  RF_CUDA(cudaGetLastError());
  return run_generic_axis<T>(c, c->field, c->field, 0, 1, c->nx, c->ny, nzh, nzh, 1, nzh, 1,
                             (long long)c->ny * nzh, (long long)c->ny * nzh, (T)1, c->stream);
}

extern "C" int rf_generate_xpass(rf_ctx* c, uint64_t seed, int potential) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  c->last_chunked = false;
  if (!c->have_power) return fail(RF_ERR_STATE, "rf_generate_xpass: call rf_set_power first");
  return c->dtype == RF_F32 ? xpass_impl<float>(c, seed, potential) : xpass_impl<double>(c, seed, potential);
}

// ------------------------------------------------------------------------------------------
// passes 2 + 3
// ------------------------------------------------------------------------------------------
template <typename T>
static int upload_table(rf_ctx* c, void* dst, const double* src, int n) {
  std::vector<T> h(n);
  for (int i = 0; i < n; ++i) h[i] = (T)src[i];
  // pageable source: the copy is staged before the call returns, ordered on the stream
  RF_CUDA(cudaMemcpyAsync(dst, h.data(), sizeof(T) * n, cudaMemcpyHostToDevice, c->stream));
  RF_CUDA(cudaStreamSynchronize(c->stream));
  return RF_OK;
}

// out = field * zs[z], used by the generic path for the light-cone factor
template <typename T>
__global__ void scale_z_kernel(T* field, long long n, int nz, const T* zs) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    field[i] *= __ldg(zs + (int)(i % nz));
}

// pass 2 (y) of local x rows [x0, x1) on the pipelined kernels (float32); source = the field buffer
// (one GPU) or the all-to-all receive buffer [src][nxl][nyl][nz/2]
static int pipe_pass2_rows(rf_ctx* c, int x0, int x1) {
  const int nzc = c->nz / 2;
  pipe::K2PParams q;
  q.tw = (const cpx<float>*)c->tw_y;
  q.rows_per_src = c->world > 1 ? c->nyl : c->ny;
  q.out = (cpx<float>*)c->field; q.out_outer = (long long)c->ny * nzc; q.out_row = nzc;
  q.z_out_base = 0;
  const void* in = c->world > 1 ? c->recv : c->field;
  const long long in_outer = (long long)(c->world > 1 ? c->nyl : c->ny) * nzc;
  const long long in_src = c->world > 1 ? (long long)c->nxl * c->nyl * nzc : in_outer * c->nxl;
  CUtensorMap min, mout;
  int e = 0;
  RF_DISPATCH_PIPE(c->ny, {
    const int tz = PipeLaunch<NN>::tile_cols();
    q.ztiles = nzc / tz;
    q.tile0 = x0 * q.ztiles;
    q.ntiles = (x1 - x0) * q.ztiles;
    q.box_rows = std::min(PipeLaunch<NN>::box_rows(), q.rows_per_src);
    const int nsrc = c->ny / q.rows_per_src;
    uint64_t din[4] = {(uint64_t)2 * nzc, (uint64_t)q.rows_per_src, (uint64_t)c->nxl, (uint64_t)nsrc};
    uint64_t sin[3] = {(uint64_t)nzc * 8, (uint64_t)in_outer * 8, (uint64_t)in_src * 8};
    uint32_t bin[4] = {(uint32_t)2 * tz, (uint32_t)q.box_rows, 1, 1};
    int e2 = make_tensor_map(&min, (void*)in, 4, din, sin, bin);
    if (e2) return e2;
    uint64_t dout[3] = {(uint64_t)2 * nzc, (uint64_t)c->ny, (uint64_t)c->nxl};
    uint64_t sout[2] = {(uint64_t)nzc * 8, (uint64_t)c->ny * nzc * 8};
    uint32_t bout[3] = {(uint32_t)2 * tz, (uint32_t)PipeLaunch<NN>::box_rows(), 1};
    e2 = make_tensor_map(&mout, c->field, 3, dout, sout, bout);
    if (e2) return e2;
    const int grid = std::min(q.ntiles, PipeLaunch<NN>::ctas_per_sm() * c->num_sms);
    e = PipeLaunch<NN>::k2p(min, mout, q, grid, c->stream);
  });
  c->launches++;
  RF_CUDA(e);
  return RF_OK;
}

// pass 3 (z, c2r) of local x rows [x0, x1) on the pipelined kernels; statistics partials (one
// (sum, sumsq) pair per CTA) are appended at c->part_used when want_stats
static int pipe_pass3_rows(rf_ctx* c, int x0, int x1, const float* zs, int want_stats) {
  pipe::K3PParams q;
  q.buf = (float*)c->field;
  q.twH = (const cpx<float>*)c->tw_zh;
  q.twN = (const cpx<float>*)c->tw_z;
  q.zscale = zs;
  q.partials = nullptr;
  int e = 0;
  RF_DISPATCH_PIPE(c->nz, {
    const int lpt = PipeLaunch<NN>::k3p_lines_per_tile();
    const long long l0 = (long long)x0 * c->ny, l1 = (long long)x1 * c->ny;
    if (l0 % lpt || l1 % lpt) return fail(RF_ERR_ARG, "pass 3: the x range does not hold whole tiles");
    q.tile0 = (int)(l0 / lpt);
    q.ntiles = (int)((l1 - l0) / lpt);
    const int grid = std::min(q.ntiles, c->num_sms);
    if (want_stats) {
      if (c->part_used + grid > c->parts_cap) return fail(RF_ERR_STATE, "pass 3: statistics buffer too small");
      q.partials = c->partials + 2 * c->part_used;
      c->part_used += grid;
    }
    e = c->k3q ? PipeLaunch<NN>::k3q(q, grid, c->stream) : PipeLaunch<NN>::k3p(q, grid, c->stream);
  });
  c->launches++;
  RF_CUDA(e);
  return RF_OK;
}

static bool pipe_pass2_ok(rf_ctx* c) {
  return c->dtype == RF_F32 && c->pipe && (c->ny <= 1024 || (c->ny <= 4096 && c->big));
}
static bool pipe_pass3_ok(rf_ctx* c) { return c->dtype == RF_F32 && c->pipe && c->nz <= 4096; }

static int finish_stats(rf_ctx* c) {
  stats_final_kernel<<<1, 256, 0, c->stream>>>(c->partials, c->nparts, c->stats_dev);
  c->launches++;
  RF_CUDA(cudaGetLastError());
  c->stats_valid = true;
  return RF_OK;
}

template <typename T>
static int finish_impl(rf_ctx* c, const double* zscale, int want_stats) {
  if (c->field_is_real) return fail(RF_ERR_STATE, "rf_finish_yz: no k-space block pending");
  const T* zs = nullptr;
  if (zscale) {
    int e = upload_table<T>(c, c->zs_dev, zscale, c->nz);
    if (e) return e;
    zs = (const T*)c->zs_dev;
  }
  if (c->fast) {
    const int nzc = c->nz / 2;
    int e = 0;
    bool done2 = false, done3 = false;
    if (c->pass2_done) {
      // pass 2 already ran per chunk (rf_chunk_pass2)
      c->pass2_done = false;
      done2 = true;
      mark(c, 2);
      mark(c, 3);
    }
    if constexpr (std::is_same<T, float>::value) {
      const int t3_per_plane = (int)((long long)c->ny * nzc * 8 / pipe::kTileBytes);
      if (!done2 && c->fuse && c->ny <= 1024 && pipe_pass2_ok(c) && t3_per_plane >= 1) {
        // y and z passes in one persistent kernel (the z pass reads the y-pass output from L2)
        pipe::K23Params f;
        pipe::K2PParams& q = f.k2;
        q.tw = (const cpx<float>*)c->tw_y;
        q.rows_per_src = c->world > 1 ? c->nyl : c->ny;
        q.out = (cpx<float>*)c->field; q.out_outer = (long long)c->ny * nzc; q.out_row = nzc;
        q.z_out_base = 0;
        const void* in = c->world > 1 ? c->recv : c->field;
        const long long in_outer = (long long)(c->world > 1 ? c->nyl : c->ny) * nzc;
        const long long in_src = c->world > 1 ? (long long)c->nxl * c->nyl * nzc : in_outer * c->nxl;
        CUtensorMap min, mout;
        RF_DISPATCH_PIPE1K(c->ny, {
          const int tz = PipeLaunch<NN>::tile_cols();
          q.ztiles = nzc / tz;
          q.tile0 = 0;
          q.ntiles = c->nxl * q.ztiles;
          q.box_rows = std::min(PipeLaunch<NN>::box_rows(), q.rows_per_src);
          const int nsrc = c->ny / q.rows_per_src;
          uint64_t din[4] = {(uint64_t)2 * nzc, (uint64_t)q.rows_per_src, (uint64_t)c->nxl, (uint64_t)nsrc};
          uint64_t sin[3] = {(uint64_t)nzc * 8, (uint64_t)in_outer * 8, (uint64_t)in_src * 8};
          uint32_t bin[4] = {(uint32_t)2 * tz, (uint32_t)q.box_rows, 1, 1};
          int e2 = make_tensor_map(&min, (void*)in, 4, din, sin, bin);
          if (e2) return e2;
          uint64_t dout[3] = {(uint64_t)2 * nzc, (uint64_t)c->ny, (uint64_t)c->nxl};
          uint64_t sout[2] = {(uint64_t)nzc * 8, (uint64_t)c->ny * nzc * 8};
          uint32_t bout[3] = {(uint32_t)2 * tz, (uint32_t)PipeLaunch<NN>::box_rows(), 1};
          e2 = make_tensor_map(&mout, c->field, 3, dout, sout, bout);
          if (e2) return e2;
          f.k3.buf = (float*)c->field;
          f.k3.twH = (const cpx<float>*)c->tw_zh;
          f.k3.twN = (const cpx<float>*)c->tw_z;
          f.k3.zscale = zs;
          f.k3.partials = nullptr;
          f.k3.ntiles = 0; f.k3.tile0 = 0;
          f.nplanes = c->nxl;
          f.t2 = q.ztiles;
          f.t3 = t3_per_plane;
          // measured best with about 5.2 rounds of tickets between a plane's y pass and its z pass
          // (lag 6 at 1024^3, 24 at 512^3: ~25 MB of y-pass output waiting in L2)
          f.lag = std::min(c->nxl, std::max(1, (21 * c->num_sms / 4) / (f.t2 + f.t3)));
          if (c->env_lag > 0) f.lag = std::min(c->nxl, c->env_lag);
          if (!c->sched) RF_CUDA(cudaMalloc((void**)&c->sched, sizeof(unsigned int) * (1 + c->nxl)));
          RF_CUDA(cudaMemsetAsync(c->sched, 0, sizeof(unsigned int) * (1 + c->nxl), c->stream));
          f.ticket = c->sched;
          f.plane_done = c->sched + 1;
          const int total = c->nxl * (f.t2 + f.t3);
          const int grid = std::min(total, c->num_sms);
          if (want_stats) {
            int e2b = ensure_partials(c, grid);
            if (e2b) return e2b;
            f.k3.partials = c->partials;
          }
          mark(c, 2);
          e = PipeLaunch<NN>::k23q(c->nz, min, mout, f, grid, c->stream);
          mark(c, 3);
          mark(c, 4);
          mark(c, 5);
        });
        c->launches++;
        RF_CUDA(e);
        done2 = done3 = true;
      }
      if (!done2 && pipe_pass2_ok(c)) {
        mark(c, 2);
        e = pipe_pass2_rows(c, 0, c->nxl);
        mark(c, 3);
        if (e) return e;
        done2 = true;
      }
    }
    if (!done2) {
      K2Params<T> p2;
      p2.tw = (const cpx<T>*)c->tw_y;
      p2.ncols = nzc;
      p2.scale = (T)1;
      p2.out = (cpx<T>*)c->field;
      p2.out_outer = (long long)c->ny * nzc;
      p2.out_row = nzc;
      if (c->world > 1) {
        p2.in = (const cpx<T>*)c->recv;
        p2.in_outer = (long long)c->nyl * nzc;
        p2.in_row = nzc;
        p2.in_src = (long long)c->nxl * c->nyl * nzc;
        p2.rows_per_src = c->nyl;
      } else {
        p2.in = (const cpx<T>*)c->field;
        p2.in_outer = (long long)c->ny * nzc;
        p2.in_row = nzc;
        p2.in_src = 0;
        p2.rows_per_src = c->ny;
      }
      RF_DISPATCH_N(c->ny, {
        int grid = c->nxl * (nzc / FftPlan<T, NN>::TZ);
        mark(c, 2);
        e = Launch<T, NN>::k2(true, p2, grid, c->stream);
        mark(c, 3);
      });
      c->launches++;
      RF_CUDA(e);
    }
    if constexpr (std::is_same<T, float>::value) {
      if (!done3 && pipe_pass3_ok(c)) {
        if (want_stats) {
          int e2 = ensure_partials(c, c->num_sms);
          if (e2) return e2;
        }
        c->part_used = 0;
        mark(c, 4);
        e = pipe_pass3_rows(c, 0, c->nxl, zs, want_stats);
        mark(c, 5);
        if (e) return e;
        if (want_stats) c->nparts = c->part_used;
        done3 = true;
      }
    }
    if (!done3) {
      K3Params<T> p3;
      p3.buf = (T*)c->field;
      p3.nlines = (long long)c->nxl * c->ny;
      p3.twH = (const cpx<T>*)c->tw_zh;
      p3.twN = (const cpx<T>*)c->tw_z;
      p3.zscale = zs;
      p3.partials = nullptr;
      RF_DISPATCH_N(c->nz, {
        int tl = Launch<T, NN>::k3_lines_per_cta();
        long long grid = (p3.nlines + tl - 1) / tl;
        if (want_stats) {
          int e2 = ensure_partials(c, grid);
          if (e2) return e2;
          p3.partials = c->partials;
        }
        mark(c, 4);
        e = Launch<T, NN>::k3(p3, (int)grid, c->stream);
        mark(c, 5);
      });
      c->launches++;
      RF_CUDA(e);
    }
  } else {
    const int nzh = c->nz / 2 + 1;
    const long long plane = (long long)c->ny * nzh;
    int e = run_generic_axis<T>(c, c->field, c->field, 0, 1, c->ny, c->nx, nzh, plane, 1, plane, 1,
                                nzh, nzh, (T)1, c->stream);
    if (e) return e;
    if (!c->recv) RF_CUDA(cudaMalloc(&c->recv, c->buf_bytes));
    // c2r out of place: complex pitch nzh -> dense real pitch nz
    e = run_generic_axis<T>(c, c->field, c->recv, 1, 1, c->nz, (long long)c->nx * c->ny, 1, nzh, 0,
                            c->nz, 0, 1, 1, (T)1, c->stream);
    if (e) return e;
    std::swap(c->field, c->recv);
    const long long n = (long long)c->nx * c->ny * c->nz;
    if (want_stats) {
      // (before the per-z factor, like the tile kernels: generate.py:219 takes the rms of delta)
      int e2 = ensure_partials(c, 1024);
      if (e2) return e2;
      stats_partial_kernel<T><<<1024, 256, 0, c->stream>>>((const T*)c->field, n, c->partials);
      c->launches++;
    }
    if (zs) {
      const int grid = (int)std::min<long long>((n + 255) / 256, 148LL * 16);
      scale_z_kernel<T><<<grid, 256, 0, c->stream>>>((T*)c->field, n, c->nz, zs);
      c->launches++;
      RF_CUDA(cudaGetLastError());
    }
  }
  if (want_stats) {
    int e = finish_stats(c);
    if (e) return e;
  }
  c->field_is_real = true;
  c->aux_is_psi = false;
  return RF_OK;
}

extern "C" int rf_finish_yz(rf_ctx* c, const double* zscale, int want_stats) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  return c->dtype == RF_F32 ? finish_impl<float>(c, zscale, want_stats)
                            : finish_impl<double>(c, zscale, want_stats);
}

// ------------------------------------------------------------------------------------------
// chunked exchange: pass 1 / copy-engine exchange / pass 2 per chunk of packed z columns
// ------------------------------------------------------------------------------------------
// timing record of one launch of a chunked step (pass 0, 1 or 2); -1 when timing is off
static int seg_begin(rf_ctx* c, int pass) {
  if (!c->timing) return -1;
  if (c->seg_used == (int)c->seg_a.size()) {
    cudaEvent_t a = nullptr, b = nullptr;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return -1;
    c->seg_a.push_back(a); c->seg_b.push_back(b); c->seg_pass.push_back(pass);
  }
  const int k = c->seg_used++;
  c->seg_pass[k] = pass;
  cudaEventRecord(c->seg_a[k], c->stream);
  return k;
}
static void seg_end(rf_ctx* c, int k) {
  if (k >= 0) cudaEventRecord(c->seg_b[k], c->stream);
}

static int pipe_tile_cols(int n) {
  RF_DISPATCH_PIPE(n, { return PipeLaunch<NN>::tile_cols(); });
  return 0;
}
static bool chunk_supported(rf_ctx* c, int nchunks) {
  if (!c->pipe || c->dtype != RF_F32 || !c->hinge_ok || c->noise) return false;
  if (c->nx > 4096 || c->ny > 4096 || c->nx < 64 || c->ny < 64) return false;
  const int nzc = c->nz / 2;
  if (nchunks < 1 || nzc % nchunks) return false;
  const int wc = nzc / nchunks;
  const int t1 = pipe_tile_cols(c->nx), t2 = pipe_tile_cols(c->ny);   // a chunk holds whole tiles
  return t1 > 0 && t2 > 0 && wc % t1 == 0 && wc % t2 == 0;
}

extern "C" int rf_chunk_config(rf_ctx* c, int nchunks) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  if (nchunks <= 0) {
    RF_CUDA(cudaStreamSynchronize(c->stream));
    if (c->send) cudaFree(c->send);
    c->send = nullptr;
    c->ck_n = 0;
    c->ck_cur = -1;
    return RF_OK;
  }
  if (!c->have_power) return fail(RF_ERR_STATE, "rf_chunk_config: call rf_set_power first");
  if (!chunk_supported(c, nchunks))
    return fail(RF_ERR_ARG, "rf_chunk_config: grid / table / chunk count not supported by the chunked path");
  if (c->nx > 1024) {
    const bool fits = c->nx == 2048 ? PipeLaunch<2048>::k1p_fits(c->lut_n) : PipeLaunch<4096>::k1p_fits(c->lut_n);
    if (!fits) return fail(RF_ERR_ARG, "rf_chunk_config: sigma table too large for the pass-1 kernel");
  }
  if (!c->send) RF_CUDA(cudaMalloc(&c->send, (size_t)c->nx * c->nyl * (c->nz / 2) * sizeof(cpx<float>)));
  if (!c->comm) RF_CUDA(cudaStreamCreateWithFlags(&c->comm, cudaStreamNonBlocking));
  if (!c->ev_k1) RF_CUDA(cudaEventCreateWithFlags(&c->ev_k1, cudaEventDisableTiming));
  if (!c->ev_x) RF_CUDA(cudaEventCreateWithFlags(&c->ev_x, cudaEventDisableTiming));
  if (!c->ev_s) RF_CUDA(cudaEventCreateWithFlags(&c->ev_s, cudaEventDisableTiming));
  c->ck_n = nchunks;
  c->ck_cur = -1;
  c->ck_reserve = 0;   // measured: no gain from keeping SMs free for the fence collectives
  if (const char* er = getenv("RF_RESERVE_SMS")) c->ck_reserve = std::min(std::max(0, atoi(er)), c->num_sms / 2);
  return RF_OK;
}

extern "C" void* rf_comm_stream(rf_ctx* c) { return c ? (void*)c->comm : nullptr; }

// comm stream waits for everything queued on the compute stream so far (start of a step: the
// previous step's pass 2 has finished reading the receive buffer once this has run)
extern "C" int rf_chunk_begin_step(rf_ctx* c) {
  if (!c || c->ck_n <= 0) return fail(RF_ERR_STATE, "rf_chunk_begin_step: chunked path not configured");
  RF_CUDA(cudaEventRecord(c->ev_s, c->stream));
  RF_CUDA(cudaStreamWaitEvent(c->comm, c->ev_s, 0));
  c->seg_used = 0;
  c->part_used = 0;
  c->zs_loaded = false;
  return RF_OK;
}

extern "C" int rf_chunk_xpass(rf_ctx* c, uint64_t seed, int potential, int chunk) {
  if (!c || c->ck_n <= 0) return fail(RF_ERR_STATE, "rf_chunk_xpass: chunked path not configured");
  if (chunk < 0 || chunk >= c->ck_n) return fail(RF_ERR_ARG, "rf_chunk_xpass: bad chunk");
  c->last_chunked = true;
  c->ck_cur = chunk;
  const int sg = seg_begin(c, 0);
  int e = xpass_impl<float>(c, seed, potential);
  seg_end(c, sg);
  c->ck_cur = -1;
  if (e) return e;
  RF_CUDA(cudaEventRecord(c->ev_k1, c->stream));
  return RF_OK;
}

// copy engines: rows x in [x0, x1) of block d of send[chunk] -> block `rank` of recv[chunk] on
// rank d (peer memory); chunk < 0: the same rows of every chunk in one pitched copy per peer
static int chunk_exchange_impl(rf_ctx* c, int chunk, int x0, int x1, bool wait_k1) {
  if (wait_k1) RF_CUDA(cudaStreamWaitEvent(c->comm, c->ev_k1, 0));
  if (c->world == 1 || x1 <= x0) return RF_OK;   // (one rank: pass 2 reads the send buffer directly)
  const size_t wc = (size_t)(c->nz / 2) / c->ck_n;
  const size_t row = (size_t)c->nyl * wc * sizeof(cpx<float>);     // one x of one block
  const size_t block = (size_t)c->nxl * row;
  const size_t chunk_off = chunk < 0 ? 0 : (size_t)chunk * c->world * block;
  for (int i = 0; i + 1 < c->world; ++i) {        // (the own block was stored in place by pass 1)
    const int d = (c->rank + 1 + i) % c->world;   // staggered: no receiver is everybody's first target
    char* peer = (char*)(c->peer_recv[d] ? c->peer_recv[d] : (d == c->rank ? c->recv : nullptr));
    if (!peer) return fail(RF_ERR_STATE, "rf_chunk_exchange: receive buffer of a peer is not mapped");
    char* dst = peer + chunk_off + (size_t)c->rank * block + (size_t)x0 * row;
    const char* src = (const char*)c->send + chunk_off + (size_t)d * block + (size_t)x0 * row;
    if (chunk >= 0)
      RF_CUDA(cudaMemcpyAsync(dst, src, (size_t)(x1 - x0) * row, cudaMemcpyDeviceToDevice, c->comm));
    else
      RF_CUDA(cudaMemcpy2DAsync(dst, (size_t)c->world * block, src, (size_t)c->world * block,
                                (size_t)(x1 - x0) * row, (size_t)c->ck_n, cudaMemcpyDeviceToDevice, c->comm));
  }
  return RF_OK;
}
extern "C" int rf_chunk_exchange(rf_ctx* c, int chunk) {
  if (!c || c->ck_n <= 0) return fail(RF_ERR_STATE, "rf_chunk_exchange: chunked path not configured");
  if (chunk < 0 || chunk >= c->ck_n) return fail(RF_ERR_ARG, "rf_chunk_exchange: bad chunk");
  return chunk_exchange_impl(c, chunk, 0, c->nxl, true);
}
extern "C" int rf_chunk_exchange_x(rf_ctx* c, int chunk, int x0, int x1, int wait_pass1) {
  if (!c || c->ck_n <= 0) return fail(RF_ERR_STATE, "rf_chunk_exchange_x: chunked path not configured");
  if (chunk >= c->ck_n || x0 < 0 || x1 > c->nxl || x0 > x1) return fail(RF_ERR_ARG, "rf_chunk_exchange_x: bad range");
  return chunk_exchange_impl(c, chunk, x0, x1, wait_pass1 != 0);
}

// call after the cross-rank fence for this chunk has been queued on the comm stream
extern "C" int rf_chunk_arrived(rf_ctx* c, int chunk) {
  if (!c || c->ck_n <= 0) return fail(RF_ERR_STATE, "rf_chunk_arrived: chunked path not configured");
  if (chunk < 0 || chunk >= 4096) return fail(RF_ERR_ARG, "rf_chunk_arrived: bad index");
  // one event per chunk (indices >= the chunk count: per x range of the split schedule): the
  // compute stream only waits for it when the pass 2 that needs it is queued
  if ((int)c->ev_xc.size() <= chunk) {
    const size_t old = c->ev_xc.size();
    c->ev_xc.resize((size_t)chunk + 1, nullptr);
    for (size_t k = old; k < c->ev_xc.size(); ++k)
      RF_CUDA(cudaEventCreateWithFlags(&c->ev_xc[k], cudaEventDisableTiming));
  }
  RF_CUDA(cudaEventRecord(c->ev_xc[chunk], c->comm));
  return RF_OK;
}

// pass 2 of rows x in [x0, x1) of one chunk (chunk < 0: of every chunk, in one launch); waits for
// event `ev` of rf_chunk_arrived first (ev >= 0)
static int chunk_pass2_impl(rf_ctx* c, int chunk, int x0, int x1, int ev) {
  if (ev >= 0) {
    if (ev >= (int)c->ev_xc.size() || !c->ev_xc[ev])
      return fail(RF_ERR_STATE, "rf_chunk_pass2: rf_chunk_arrived has not been called for this index");
    RF_CUDA(cudaStreamWaitEvent(c->stream, c->ev_xc[ev], 0));
  }
  if (x1 <= x0) return RF_OK;
  const int nzc = c->nz / 2, wc = nzc / c->ck_n;
  const int nck = chunk < 0 ? c->ck_n : 1;
  pipe::K2PParams q;
  q.tw = (const cpx<float>*)c->tw_y;
  q.out = (cpx<float>*)c->field;
  q.out_outer = (long long)c->ny * nzc;
  q.out_row = nzc;
  q.z_out_base = chunk < 0 ? 0 : chunk * wc;
  const size_t chunk_elems = (size_t)c->nx * c->nyl * wc;
  const cpx<float>* in = (const cpx<float>*)(c->world > 1 ? c->recv : c->send) +
                         (size_t)(chunk < 0 ? 0 : chunk) * chunk_elems;
  q.rows_per_src = c->world > 1 ? c->nyl : c->ny;
  int e = 0;
  CUtensorMap min, mout;
  RF_DISPATCH_PIPE(c->ny, {
    const int tz = PipeLaunch<NN>::tile_cols();
    q.ztiles = wc / tz;
    q.tile0 = x0 * q.ztiles;
    q.ck_tiles = (x1 - x0) * q.ztiles;
    q.ck_wc = wc;
    q.ntiles = nck * q.ck_tiles;
    q.box_rows = std::min(PipeLaunch<NN>::box_rows(), q.rows_per_src);
    const int nsrc = c->ny / q.rows_per_src;
    q.nsrc = nsrc;
    // [chunk][src][nxl][rows_per_src][wc] (single GPU: [chunk][nx][ny][wc]); a chunk's source
    // blocks follow the previous chunk's, so (chunk, src) is one dimension of the map
    uint64_t din[4] = {(uint64_t)2 * wc, (uint64_t)q.rows_per_src, (uint64_t)c->nxl, (uint64_t)nsrc * nck};
    uint64_t sin[3] = {(uint64_t)wc * 8, (uint64_t)q.rows_per_src * wc * 8,
                       (uint64_t)c->nxl * q.rows_per_src * wc * 8};
    uint32_t bin[4] = {(uint32_t)2 * tz, (uint32_t)q.box_rows, 1, 1};
    int e2 = make_tensor_map(&min, (void*)in, 4, din, sin, bin);
    if (e2) return e2;
    uint64_t dout[3] = {(uint64_t)2 * nzc, (uint64_t)c->ny, (uint64_t)c->nxl};
    uint64_t sout[2] = {(uint64_t)nzc * 8, (uint64_t)c->ny * nzc * 8};
    uint32_t bout[3] = {(uint32_t)2 * tz, (uint32_t)PipeLaunch<NN>::box_rows(), 1};
    e2 = make_tensor_map(&mout, c->field, 3, dout, sout, bout);
    if (e2) return e2;
    const int grid = std::min(q.ntiles, PipeLaunch<NN>::ctas_per_sm() * (c->num_sms - c->ck_reserve));
    const int sg = seg_begin(c, 1);
    e = PipeLaunch<NN>::k2p(min, mout, q, grid, c->stream);
    seg_end(c, sg);
  });
  c->launches++;
  RF_CUDA(e);
  return RF_OK;
}
extern "C" int rf_chunk_pass2(rf_ctx* c, int chunk) {
  if (!c || c->ck_n <= 0) return fail(RF_ERR_STATE, "rf_chunk_pass2: chunked path not configured");
  if (chunk < 0 || chunk >= c->ck_n) return fail(RF_ERR_ARG, "rf_chunk_pass2: bad chunk");
  int e = chunk_pass2_impl(c, chunk, 0, c->nxl, chunk);
  if (e) return e;
  if (chunk == c->ck_n - 1) c->pass2_done = true;   // rf_finish_yz now runs pass 3 only
  return RF_OK;
}
extern "C" int rf_chunk_pass2_x(rf_ctx* c, int chunk, int x0, int x1, int wait_index) {
  if (!c || c->ck_n <= 0) return fail(RF_ERR_STATE, "rf_chunk_pass2_x: chunked path not configured");
  if (chunk >= c->ck_n || x0 < 0 || x1 > c->nxl || x0 > x1)
    return fail(RF_ERR_ARG, "rf_chunk_pass2_x: bad chunk or range");
  return chunk_pass2_impl(c, chunk, x0, x1, wait_index);
}

// pass 3 of rows x in [x0, x1) once pass 2 of every chunk has been queued for them; `last`
// closes the step (statistics over all ranges, the field becomes the real-space field)
extern "C" int rf_chunk_pass3_x(rf_ctx* c, int x0, int x1, const double* zscale, int want_stats, int last) {
  if (!c || c->ck_n <= 0) return fail(RF_ERR_STATE, "rf_chunk_pass3_x: chunked path not configured");
  if (x0 < 0 || x1 > c->nxl || x0 > x1) return fail(RF_ERR_ARG, "rf_chunk_pass3_x: bad range");
  if (!pipe_pass3_ok(c)) return fail(RF_ERR_STATE, "rf_chunk_pass3_x: pipelined pass 3 not available");
  const float* zs = nullptr;
  if (zscale) {
    if (!c->zs_loaded) {
      int e = upload_table<float>(c, c->zs_dev, zscale, c->nz);
      if (e) return e;
      c->zs_loaded = true;
    }
    zs = (const float*)c->zs_dev;
  }
  if (want_stats && c->parts_cap < (long long)c->num_sms * 64) {
    if (c->part_used) return fail(RF_ERR_STATE, "rf_chunk_pass3_x: partials buffer too small");
    int e = ensure_partials(c, (long long)c->num_sms * 64);
    if (e) return e;
  }
  if (x1 > x0) {
    const int sg = seg_begin(c, 2);
    int e = pipe_pass3_rows(c, x0, x1, zs, want_stats);
    seg_end(c, sg);
    if (e) return e;
  }
  if (last) {
    if (want_stats) {
      c->nparts = c->part_used;
      int e = finish_stats(c);
      if (e) return e;
    }
    c->pass2_done = false;
    c->field_is_real = true;
    c->aux_is_psi = false;
  }
  return RF_OK;
}

// ------------------------------------------------------------------------------------------
// streamed output: pass 2 + pass 3 per range of x rows, each range's
// device -> host copy on the copy stream while the next range is transformed
// ------------------------------------------------------------------------------------------
static int ensure_copy_stream(rf_ctx* c) {
  if (!c->copy_stream) RF_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
  if (!c->ev_copy) RF_CUDA(cudaEventCreateWithFlags(&c->ev_copy, cudaEventDisableTiming));
  return RF_OK;
}

extern "C" int rf_copy_rows(rf_ctx* c, int which, int x0, int x1, void* host_dst, int async) {
  if (!c || !host_dst) return fail(RF_ERR_ARG, "rf_copy_rows: null argument");
  if (x0 < 0 || x1 > c->nxl || x0 > x1) return fail(RF_ERR_ARG, "rf_copy_rows: bad x range");
  const char* src = nullptr;
  if (which == RF_BUF_FIELD) {
    if (!c->field_is_real) return fail(RF_ERR_STATE, "rf_copy_rows: no real-space field");
    src = (const char*)c->field;
  } else if (which == RF_BUF_AUX) {
    if (!c->aux_is_psi) return fail(RF_ERR_STATE, "rf_copy_rows: no lensing potential in the auxiliary buffer");
    src = (const char*)c->recv;
  } else {
    return fail(RF_ERR_ARG, "rf_copy_rows: which must be RF_BUF_FIELD or RF_BUF_AUX");
  }
  int e = ensure_copy_stream(c);
  if (e) return e;
  const size_t row = (size_t)c->ny * c->nz * c->esize;
  // after everything queued on the compute stream so far
  RF_CUDA(cudaEventRecord(c->ev_copy, c->stream));
  RF_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_copy, 0));
  if (x1 > x0)
    RF_CUDA(cudaMemcpyAsync(host_dst, src + (size_t)x0 * row, (size_t)(x1 - x0) * row, cudaMemcpyDeviceToHost,
                            c->copy_stream));
  if (!async) RF_CUDA(cudaStreamSynchronize(c->copy_stream));
  return RF_OK;
}

extern "C" int rf_copy_wait(rf_ctx* c) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  if (c->copy_stream) RF_CUDA(cudaStreamSynchronize(c->copy_stream));
  return RF_OK;
}

extern "C" int rf_finish_yz_streamed(rf_ctx* c, const double* zscale, int want_stats, void* host_dst,
                                     int nranges) {
  if (!c || !host_dst) return fail(RF_ERR_ARG, "rf_finish_yz_streamed: null argument");
  if (c->field_is_real) return fail(RF_ERR_STATE, "rf_finish_yz_streamed: no k-space block pending");
  int e = ensure_copy_stream(c);
  if (e) return e;
  const bool ranged = c->fast && !c->pass2_done && pipe_pass2_ok(c) && pipe_pass3_ok(c) && nranges > 1;
  int align = 1;
  if (ranged) {
    // ranges hold whole pass-3 tiles
    int lpt = 1;
    RF_DISPATCH_PIPE(c->nz, { lpt = PipeLaunch<NN>::k3p_lines_per_tile(); });
    int g = lpt, y = c->ny;
    while (y) { int t = g % y; g = y; y = t; }
    align = lpt / g;
  }
  const int units = (ranged && c->nxl % align == 0) ? c->nxl / align : 0;
  const size_t row = (size_t)c->ny * c->nz * c->esize;
  if (units < 2) {
    // one range: the whole-field kernels (fused y + z where selected), then one copy
    e = rf_finish_yz(c, zscale, want_stats);
    if (e) return e;
    return rf_copy_rows(c, RF_BUF_FIELD, 0, c->nxl, host_dst, 1);
  }
  nranges = std::min(nranges, units);
  const float* zs = nullptr;
  if (zscale) {
    e = upload_table<float>(c, c->zs_dev, zscale, c->nz);
    if (e) return e;
    zs = (const float*)c->zs_dev;
  }
  if (want_stats) {
    e = ensure_partials(c, (long long)c->num_sms * nranges);
    if (e) return e;
  }
  c->part_used = 0;
  while ((int)c->ev_rng.size() < nranges) {
    cudaEvent_t ev = nullptr;
    RF_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    c->ev_rng.push_back(ev);
  }
  mark(c, 2);
  for (int r = 0; r < nranges; ++r) {
    const int x0 = align * (int)((long long)units * r / nranges), x1 = align * (int)((long long)units * (r + 1) / nranges);
    if (x1 <= x0) continue;
    if ((e = pipe_pass2_rows(c, x0, x1))) return e;
    if ((e = pipe_pass3_rows(c, x0, x1, zs, want_stats))) return e;
    RF_CUDA(cudaEventRecord(c->ev_rng[r], c->stream));
    RF_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_rng[r], 0));
    RF_CUDA(cudaMemcpyAsync((char*)host_dst + (size_t)x0 * row, (const char*)c->field + (size_t)x0 * row,
                            (size_t)(x1 - x0) * row, cudaMemcpyDeviceToHost, c->copy_stream));
  }
  mark(c, 3); mark(c, 4); mark(c, 5);
  if (want_stats) {
    c->nparts = c->part_used;
    if ((e = finish_stats(c))) return e;
  }
  c->field_is_real = true;
  c->aux_is_psi = false;
  return RF_OK;
}

extern "C" int rf_generate_delta(rf_ctx* c, uint64_t seed, int potential, const double* zscale,
                                 int want_stats) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  if (c->world != 1) return fail(RF_ERR_STATE, "rf_generate_delta: single-GPU only; use xpass/finish");
  int e = rf_generate_xpass(c, seed, potential);
  if (e) return e;
  return rf_finish_yz(c, zscale, want_stats);
}

extern "C" int rf_stats(rf_ctx* c, double* sum, double* sumsq, long long* count) {
  if (!c) return fail(RF_ERR_ARG, "null ctx");
  if (!c->stats_valid) return fail(RF_ERR_STATE, "rf_stats: no statistics were accumulated");
  double h[2];
  RF_CUDA(cudaMemcpyAsync(h, c->stats_dev, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
  RF_CUDA(cudaStreamSynchronize(c->stream));
  if (sum) *sum = h[0];
  if (sumsq) *sumsq = h[1];
  if (count) *count = rf_field_elems(c);
  return RF_OK;
}

// ------------------------------------------------------------------------------------------
// density epilogue
// ------------------------------------------------------------------------------------------
template <typename T>
static int density_impl(rf_ctx* c, int lognormal, double sigma, const double* growth, const double* rho) {
  std::vector<double> a(c->nz), b(c->nz);
  for (int z = 0; z < c->nz; ++z) {
    if (lognormal) {
      double t = 1.0 + (sigma * growth[z]) * (sigma * growth[z]);
      a[z] = std::sqrt(std::log(t)) / sigma;
      b[z] = rho[z] / std::sqrt(t);
    } else {
      a[z] = growth[z];
      b[z] = rho[z];
    }
  }
  int e = upload_table<T>(c, c->da_dev, a.data(), c->nz);
  if (e) return e;
  e = upload_table<T>(c, c->db_dev, b.data(), c->nz);
  if (e) return e;
  const long long n = rf_field_elems(c);
  if (c->nz % 4 == 0) {
    const long long n4 = n / 4;
    int grid = (int)std::min<long long>((n4 + 255) / 256, 148LL * 32);
    density_kernel_v4<T><<<grid, 256, 0, c->stream>>>((T*)c->field, n4, c->nz, (const T*)c->da_dev,
                                                      (const T*)c->db_dev, lognormal);
  } else {
    int grid = (int)std::min<long long>((n + 255) / 256, 148LL * 32);
    density_kernel<T><<<grid, 256, 0, c->stream>>>((T*)c->field, n, c->nz, (const T*)c->da_dev,
                                                   (const T*)c->db_dev, lognormal);
  }
  c->launches++;
  RF_CUDA(cudaGetLastError());
  return RF_OK;
}

extern "C" int rf_density(rf_ctx* c, int lognormal, double sigma, const double* growth, const double* rho) {
  if (!c || !growth || !rho) return fail(RF_ERR_ARG, "rf_density: null argument");
  if (!c->field_is_real) return fail(RF_ERR_STATE, "rf_density: no real-space field");
  if (lognormal && !(sigma > 0)) return fail(RF_ERR_ARG, "rf_density: sigma must be > 0");
  return c->dtype == RF_F32 ? density_impl<float>(c, lognormal, sigma, growth, rho)
                            : density_impl<double>(c, lognormal, sigma, growth, rho);
}

// ------------------------------------------------------------------------------------------
// lensing potential along the line of sight (generate.py:353-418), in place on the real field
// ------------------------------------------------------------------------------------------
template <typename T>
static int lensing_impl(rf_ctx* c, int i_min, const double* tables5) {
  const int nz = c->nz;
  if (!c->lens_dev) RF_CUDA(cudaMalloc((void**)&c->lens_dev, sizeof(double) * 5 * nz));
  // pageable source: staged before the call returns, ordered on the stream
  RF_CUDA(cudaMemcpyAsync(c->lens_dev, tables5, sizeof(double) * 5 * nz, cudaMemcpyHostToDevice, c->stream));
  RF_CUDA(cudaStreamSynchronize(c->stream));
  LensTables tb;
  tb.cotk = c->lens_dev; tb.pa = c->lens_dev + nz; tb.pb = c->lens_dev + 2 * nz;
  tb.pd = c->lens_dev + 3 * nz; tb.pt = c->lens_dev + 4 * nz;
  if (!c->recv) RF_CUDA(cudaMalloc(&c->recv, c->buf_bytes));
  T* const psi = (T*)c->recv;   // generate.py:398-399: psi is a separate array, dPhi stays intact
  const long long nlines4 = (long long)c->nxl * c->ny;
  const bool lens4 = c->lens4;
  constexpr int S8 = LensS<T>::value;
  auto go = [&](auto kern, int S, int tpb) {
    const int per_sm = std::max(1, (sizeof(T) == 4 && S == 4 ? 768 : 512) / tpb);
    const int lpc = tpb / (nz / S);
    const int grid = (int)std::min<long long>((nlines4 + lpc - 1) / lpc, (long long)c->num_sms * per_sm);
    kern<<<grid, tpb, 0, c->stream>>>((const T*)c->field, psi, nlines4, nz, i_min, tb);
    c->launches++;
    c->aux_is_psi = true;
    return cudaGetLastError();
  };
  // fixed sample ownership: coefficients in registers, 16-byte vectors per thread and line
  // one line group per CTA where a line takes 128 threads or more (the per-line barrier then
  // spans as few warps as possible), several lines per 128-thread CTA below that
  auto pick = [&](auto s_tag) {
    constexpr int S = decltype(s_tag)::value;
    const int G = nz / S;
    return G <= 128 ? go(lensing_scan4_kernel<T, 128, S>, S, 128)
         : G <= 256 ? go(lensing_scan4_kernel<T, 256, S>, S, 256)
                    : go(lensing_scan4_kernel<T, 512, S>, S, 512);
  };
  if (lens4 && lens4_supported(nz, S8)) {
    RF_CUDA(pick(std::integral_constant<int, S8>()));
    return RF_OK;
  }
  if (lens4 && S8 != 4 && lens4_supported(nz, 4)) {
    RF_CUDA(pick(std::integral_constant<int, 4>()));
    return RF_OK;
  }
  int warps = 16;
  while (warps > 1 && lens_smem_bytes(nz, warps, sizeof(T)) > 112 * 1024) warps >>= 1;
  const size_t smem = lens_smem_bytes(nz, warps, sizeof(T));
  if (smem > 227 * 1024) return fail(RF_ERR_ARG, "rf_lensing: nz too large");
  RF_CUDA(cudaFuncSetAttribute(lensing_scan_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long nlines = (long long)c->nxl * c->ny;
  const int per_sm = std::max(1, (int)((size_t)228 * 1024 / (smem + 1024)));
  const int grid = (int)std::min<long long>((nlines + warps - 1) / warps, (long long)c->num_sms * per_sm);
  lensing_scan_kernel<T><<<grid, 32 * warps, smem, c->stream>>>((const T*)c->field, psi, nlines, nz, i_min, tb);
  c->launches++;
  c->aux_is_psi = true;
  RF_CUDA(cudaGetLastError());
  return RF_OK;
}

extern "C" int rf_lensing(rf_ctx* c, int i_min, const double* tables5) {
  if (!c || !tables5) return fail(RF_ERR_ARG, "rf_lensing: null argument");
  if (!c->field_is_real) return fail(RF_ERR_STATE, "rf_lensing: no real-space field");
  if (i_min < 0 || i_min >= c->nz) return fail(RF_ERR_ARG, "rf_lensing: i_min out of range");
  return c->dtype == RF_F32 ? lensing_impl<float>(c, i_min, tables5) : lensing_impl<double>(c, i_min, tables5);
}

extern "C" int rf_copy_field(rf_ctx* c, void* host_dst) {
  if (!c || !host_dst) return fail(RF_ERR_ARG, "rf_copy_field: null argument");
  if (!c->field_is_real) return fail(RF_ERR_STATE, "rf_copy_field: no real-space field");
  RF_CUDA(cudaMemcpyAsync(host_dst, c->field, (size_t)rf_field_elems(c) * c->esize,
                          cudaMemcpyDeviceToHost, c->stream));
  RF_CUDA(cudaStreamSynchronize(c->stream));
  return RF_OK;
}

// ------------------------------------------------------------------------------------------
// stage-level k-space export
// ------------------------------------------------------------------------------------------
template <typename T>
static int kspace_impl(rf_ctx* c, uint64_t seed, int potential, void* host, int32_t* host_bins) {
  ModeGeom<T> g = make_geom<T>(c, seed, potential);
  SigmaTable<T> tab = make_tab<T>(c);
  const int nzh = c->nz / 2 + 1;
  const long long total = (long long)c->nx * c->nyl * nzh;
  // generated into scratch memory: in the reference, looking at plan_c2r.data_in / generator.potential
  // has no side effect on the field (generate.py:84,200-217)
  cpx<T>* scratch = nullptr;
  int32_t* dbins = nullptr;
  RF_CUDA(cudaMalloc((void**)&scratch, sizeof(cpx<T>) * total));
  if (host_bins && cudaMalloc((void**)&dbins, sizeof(int32_t) * total) != cudaSuccess) {
    cudaFree(scratch);
    return fail(RF_ERR_CUDA, "rf_kspace_public: out of device memory");
  }
  const double N3 = (double)c->nx * c->ny * c->nz;
  int grid = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
  gen_kspace_public_kernel<T><<<grid, 256, 0, c->stream>>>(g, tab, scratch, dbins, (T)N3, c->rank * c->nyl, c->nyl);
  c->launches++;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(host, scratch, sizeof(cpx<T>) * total, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess && host_bins)
    e = cudaMemcpyAsync(host_bins, dbins, sizeof(int32_t) * total, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cudaFree(scratch);
  if (dbins) cudaFree(dbins);
  RF_CUDA(e);
  return RF_OK;
}

extern "C" int rf_kspace_public(rf_ctx* c, uint64_t seed, int potential, void* host_cplx, int32_t* host_bins) {
  if (!c || !host_cplx) return fail(RF_ERR_ARG, "rf_kspace_public: null argument");
  if (!c->have_power) return fail(RF_ERR_STATE, "rf_kspace_public: call rf_set_power first");
  return c->dtype == RF_F32 ? kspace_impl<float>(c, seed, potential, host_cplx, host_bins)
                            : kspace_impl<double>(c, seed, potential, host_cplx, host_bins);
}

// ------------------------------------------------------------------------------------------
// P(k) estimator: forward r2c of the real field (the reference's unused
// plan_r2c seam, generate.py:79-80, transform.py:278-301) + shell binning on the device
// ------------------------------------------------------------------------------------------
// |F(k)|^2 of every mode of the half-complex array, binned by |k| into nbins shells.  Modes of the
// interior planes 0 < iz < nz/2 stand for themselves and their Hermitian partner (weight 2), the
// planes iz in {0, nz/2} hold both partners (weight 1): sums and counts are over the full k-space.
template <typename T>
__global__ void pk_bin_kernel(const cpx<T>* f, int nx, int ny, int nz, double k0x, double k0y, double k0z,
                              int nbins, const double* __restrict__ edges, double* __restrict__ sums,
                              unsigned long long* __restrict__ counts) {
  extern __shared__ __align__(16) unsigned char pk_smem[];
  double* ss = reinterpret_cast<double*>(pk_smem);
  unsigned long long* sc = reinterpret_cast<unsigned long long*>(ss + nbins);
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) { ss[b] = 0.0; sc[b] = 0ull; }
  __syncthreads();
  const int nzh = nz / 2 + 1;
  const long long total = (long long)nx * ny * nzh;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int iz = (int)(i % nzh);
    const long long r = i / nzh;
    const int iy = (int)(r % ny), ix = (int)(r / ny);
    const double jx = ix <= nx / 2 ? ix : ix - nx, jy = iy <= ny / 2 ? iy : iy - ny, jz = iz;
    const double k = sqrt(jx * jx * k0x * k0x + jy * jy * k0y * k0y + jz * jz * k0z * k0z);
    if (!(k >= edges[0]) || k >= edges[nbins]) continue;
    int lo = 0, hi = nbins;   // edges[lo] <= k < edges[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (edges[mid] <= k) lo = mid; else hi = mid;
    }
    const cpx<T> v = f[i];
    const double p = (double)v.x * (double)v.x + (double)v.y * (double)v.y;
    const int w = (iz == 0 || iz == nz / 2) ? 1 : 2;
    atomicAdd(&ss[lo], w * p);
    atomicAdd(&sc[lo], (unsigned long long)w);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) {
    if (sc[b]) {
      atomicAdd(&sums[b], ss[b]);
      atomicAdd(&counts[b], sc[b]);
    }
  }
}

// |G|^2 of the packed-Nyquist layout [nx][ny][nz/2] the fast forward transform leaves (rf_k3f.cuh): column
// 0 holds C = F0 + i FN, the transforms of the two real planes iz = 0 and iz = nz/2, which come apart
// as F0[k] = (C[k] + conj C[-k]) / 2, FN[k] = (C[k] - conj C[-k]) / 2i with -k = (-ix, -iy).
__global__ void pk_bin_packed_kernel(const cpx<float>* f, int nx, int ny, int nz, double k0x, double k0y, double k0z,
                                     int nbins, const double* __restrict__ edges, double* __restrict__ sums,
                                     unsigned long long* __restrict__ counts) {
  extern __shared__ __align__(16) unsigned char pk_smem[];
  double* ss = reinterpret_cast<double*>(pk_smem);
  unsigned long long* sc = reinterpret_cast<unsigned long long*>(ss + nbins);
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) { ss[b] = 0.0; sc[b] = 0ull; }
  __syncthreads();
  const int nzc = nz / 2;
  const long long total = (long long)nx * ny * nzc;
  auto bin_of = [&](double k) {
    if (!(k >= edges[0]) || k >= edges[nbins]) return -1;
    int lo = 0, hi = nbins;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (edges[mid] <= k) lo = mid; else hi = mid;
    }
    return lo;
  };
  // The lanes of a warp hold consecutive iz of one line, so their shells are one or two values:
  // one shuffle reduction and one shared-memory atomic per distinct shell and warp instead of 32
  // contended atomics (1024^3: 20 ms -> a few ms).
  auto add_warp = [&](int bin, double p, int w) {
    unsigned todo = __ballot_sync(0xffffffffu, bin >= 0);
    while (todo) {
      const int b = __shfl_sync(0xffffffffu, bin, __ffs(todo) - 1);
      const bool mine = bin == b;
      double v = mine ? w * p : 0.0;
      int n = mine ? w : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        v += __shfl_xor_sync(0xffffffffu, v, o);
        n += __shfl_xor_sync(0xffffffffu, n, o);
      }
      if ((threadIdx.x & 31) == 0) {
        atomicAdd(&ss[b], v);
        atomicAdd(&sc[b], (unsigned long long)n);
      }
      todo &= ~__ballot_sync(0xffffffffu, mine);
    }
  };
  // whole warps walk the array together (the loop bound is rounded up to a multiple of 32)
  const long long total32 = (total + 31) / 32 * 32;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total32;
       i += (long long)gridDim.x * blockDim.x) {
    int bin_a = -1, bin_b = -1;
    double pa = 0.0, pb = 0.0;
    int wa = 0;
    if (i < total) {
      const int iz = (int)(i % nzc);
      const long long r = i / nzc;
      const int iy = (int)(r % ny), ix = (int)(r / ny);
      const double jx = ix <= nx / 2 ? ix : ix - nx, jy = iy <= ny / 2 ? iy : iy - ny;
      const double kxy2 = jx * jx * k0x * k0x + jy * jy * k0y * k0y;
      const cpx<float> v = f[i];
      if (iz > 0) {
        bin_a = bin_of(sqrt(kxy2 + (double)iz * iz * k0z * k0z));
        pa = (double)v.x * v.x + (double)v.y * v.y;
        wa = 2;
      } else {
        const cpx<float> m = f[((long long)((nx - ix) % nx) * ny + (ny - iy) % ny) * nzc];
        const double ar = 0.5 * ((double)v.x + m.x), ai = 0.5 * ((double)v.y - m.y);     // F0
        const double br = 0.5 * ((double)v.y + m.y), bi = -0.5 * ((double)v.x - m.x);    // FN
        bin_a = bin_of(sqrt(kxy2));
        pa = ar * ar + ai * ai;
        wa = 1;
        bin_b = bin_of(sqrt(kxy2 + (double)nzc * nzc * k0z * k0z));
        pb = br * br + bi * bi;
      }
    }
    add_warp(bin_a, pa, wa);
    if (__any_sync(0xffffffffu, bin_b >= 0)) add_warp(bin_b, pb, 1);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) {
    if (sc[b]) {
      atomicAdd(&sums[b], ss[b]);
      atomicAdd(&counts[b], sc[b]);
    }
  }
}

// in-place inverse-sign FFT along a strided axis of a packed [..][..][nzc] complex buffer on the pipelined
// y-pass kernels: lines of n points `row_stride` elements apart, `nouter` of them `outer_stride` apart
static int pipe_strided_pass(rf_ctx* c, void* buf, int n, const void* tw, long long row_stride, int nouter,
                             long long outer_stride) {
  const int nzc = c->nz / 2;
  pipe::K2PParams q;
  q.tw = (const cpx<float>*)tw;
  q.rows_per_src = n;
  q.out = (cpx<float>*)buf; q.out_outer = outer_stride; q.out_row = row_stride;
  q.z_out_base = 0;
  CUtensorMap min, mout;
  int e = 0;
  RF_DISPATCH_PIPE(n, {
    const int tz = PipeLaunch<NN>::tile_cols();
    if (nzc % tz) return fail(RF_ERR_ARG, "strided pass: nz/2 is not a multiple of the tile width");
    q.ztiles = nzc / tz;
    q.tile0 = 0;
    q.ntiles = nouter * q.ztiles;
    q.box_rows = std::min(PipeLaunch<NN>::box_rows(), n);
    uint64_t din[4] = {(uint64_t)2 * nzc, (uint64_t)n, (uint64_t)nouter, 1};
    uint64_t sin[3] = {(uint64_t)row_stride * 8, (uint64_t)outer_stride * 8, (uint64_t)row_stride * 8 * n};
    uint32_t bin[4] = {(uint32_t)2 * tz, (uint32_t)q.box_rows, 1, 1};
    int e2 = make_tensor_map(&min, buf, 4, din, sin, bin);
    if (e2) return e2;
    uint64_t dout[3] = {(uint64_t)2 * nzc, (uint64_t)n, (uint64_t)nouter};
    uint64_t sout[2] = {(uint64_t)row_stride * 8, (uint64_t)outer_stride * 8};
    uint32_t bout[3] = {(uint32_t)2 * tz, (uint32_t)PipeLaunch<NN>::box_rows(), 1};
    e2 = make_tensor_map(&mout, buf, 3, dout, sout, bout);
    if (e2) return e2;
    const int grid = std::min(q.ntiles, PipeLaunch<NN>::ctas_per_sm() * c->num_sms);
    e = PipeLaunch<NN>::k2p(min, mout, q, grid, c->stream);
  });
  c->launches++;
  RF_CUDA(e);
  return RF_OK;
}

// forward transform on the fast kernels: real field -> copy in the auxiliary buffer -> r2c along z (k3f),
// then the inverse-sign y and x passes (|F| = |conj F|); returns with the packed spectrum in c->recv
static int r2c_fast(rf_ctx* c) {
  const size_t bytes = (size_t)c->nx * c->ny * c->nz * sizeof(float);
  RF_CUDA(cudaMemcpyAsync(c->recv, c->field, bytes, cudaMemcpyDeviceToDevice, c->stream));
  c->launches++;
  pipe::K3PParams q;
  q.buf = (float*)c->recv;
  q.twH = (const cpx<float>*)c->tw_zh;
  q.twN = (const cpx<float>*)c->tw_z;
  q.zscale = nullptr;
  q.partials = nullptr;
  q.tile0 = 0;
  int e = 0;
  RF_DISPATCH_PIPE(c->nz, {
    q.ntiles = (int)((long long)c->nx * c->ny / PipeLaunch<NN>::k3p_lines_per_tile());
    e = PipeLaunch<NN>::k3f(q, std::min(q.ntiles, c->num_sms), c->stream);
  });
  c->launches++;
  RF_CUDA(e);
  const long long nzc = c->nz / 2;
  e = pipe_strided_pass(c, c->recv, c->ny, c->tw_y, nzc, c->nx, (long long)c->ny * nzc);   // y
  if (e) return e;
  return pipe_strided_pass(c, c->recv, c->nx, c->tw_x, (long long)c->ny * nzc, c->ny, nzc);   // x
}

template <typename T>
static int r2c_power_impl(rf_ctx* c, int nbins, const double* edges, double* out_pk, long long* out_counts) {
  const int nx = c->nx, ny = c->ny, nz = c->nz, nzh = nz / 2 + 1;
  // forward transform out of place: field (real, pitch nz) -> auxiliary buffer (complex, pitch nz/2+1);
  // the real field stays as it is
  if (!c->recv) RF_CUDA(cudaMalloc(&c->recv, c->buf_bytes));
  c->aux_is_psi = false;
  // float32 grids of the tile kernels (every axis a power of two in [64, 4096], nz >= 128): the pipelined
  // kernels of the hot path, run in the forward direction; everything else: any-length line transforms
  bool fast = false;
  if constexpr (std::is_same<T, float>::value)
    fast = c->pipe && c->nz >= 128 && (c->ny <= 1024 || c->big) && (c->nx <= 1024 || c->big) &&
           (long long)c->nx * c->ny % 256 == 0;
  int e = 0;
  if (fast) {
    e = r2c_fast(c);
    if (e) return e;
  } else {
    const long long plane = (long long)ny * nzh;
    e = run_generic_axis<T>(c, c->field, c->recv, 2, 0, nz, (long long)nx * ny, 1, nz, 0, nzh, 0, 1, 1, (T)1, c->stream);
    if (e) return e;
    e = run_generic_axis<T>(c, c->recv, c->recv, 0, 0, ny, nx, nzh, plane, 1, plane, 1, nzh, nzh, (T)1, c->stream);
    if (e) return e;
    e = run_generic_axis<T>(c, c->recv, c->recv, 0, 0, nx, ny, nzh, nzh, 1, nzh, 1, plane, plane, (T)1, c->stream);
    if (e) return e;
  }
  double* dev = nullptr;
  const size_t bytes = sizeof(double) * (nbins + 1) + sizeof(double) * nbins + sizeof(unsigned long long) * nbins;
  RF_CUDA(cudaMalloc((void**)&dev, bytes));
  double* d_edges = dev;
  double* d_sums = dev + nbins + 1;
  unsigned long long* d_counts = reinterpret_cast<unsigned long long*>(d_sums + nbins);
  cudaError_t ce = cudaMemcpyAsync(d_edges, edges, sizeof(double) * (nbins + 1), cudaMemcpyHostToDevice, c->stream);
  if (ce == cudaSuccess) ce = cudaMemsetAsync(d_sums, 0, bytes - sizeof(double) * (nbins + 1), c->stream);
  const double two_pi = 6.283185307179586476925286766559;
  const long long total = (long long)nx * ny * nzh;
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)c->num_sms * 8);
  if (ce == cudaSuccess) {
    if (fast)
      pk_bin_packed_kernel<<<grid, 256, nbins * 16, c->stream>>>((const cpx<float>*)c->recv, nx, ny, nz,
                                                                 two_pi / (nx * c->spacing), two_pi / (ny * c->spacing),
                                                                 two_pi / (nz * c->spacing), nbins, d_edges, d_sums, d_counts);
    else
      pk_bin_kernel<T><<<grid, 256, nbins * 16, c->stream>>>((const cpx<T>*)c->recv, nx, ny, nz, two_pi / (nx * c->spacing),
                                                          two_pi / (ny * c->spacing), two_pi / (nz * c->spacing), nbins,
                                                          d_edges, d_sums, d_counts);
    c->launches++;
    ce = cudaGetLastError();
  }
  std::vector<double> sums(nbins);
  std::vector<unsigned long long> counts(nbins);
  if (ce == cudaSuccess) ce = cudaMemcpyAsync(sums.data(), d_sums, sizeof(double) * nbins, cudaMemcpyDeviceToHost, c->stream);
  if (ce == cudaSuccess)
    ce = cudaMemcpyAsync(counts.data(), d_counts, sizeof(unsigned long long) * nbins, cudaMemcpyDeviceToHost, c->stream);
  if (ce == cudaSuccess) ce = cudaStreamSynchronize(c->stream);
  cudaFree(dev);
  RF_CUDA(ce);
  // P(k) = <|F(k)|^2> V / N3^2 with F the unnormalised forward transform (numpy rfftn convention)
  const double N3 = (double)nx * ny * nz, V = N3 * c->spacing * c->spacing * c->spacing;
  for (int b = 0; b < nbins; ++b) {
    out_pk[b] = counts[b] ? sums[b] / (double)counts[b] * V / (N3 * N3) : 0.0;
    if (out_counts) out_counts[b] = (long long)counts[b];
  }
  return RF_OK;
}

extern "C" int rf_r2c_power(rf_ctx* c, int nbins, const double* k_edges, double* out_pk, long long* out_counts) {
  if (!c || !k_edges || !out_pk) return fail(RF_ERR_ARG, "rf_r2c_power: null argument");
  if (nbins < 1 || nbins > 2048) return fail(RF_ERR_ARG, "rf_r2c_power: need 1 <= nbins <= 2048");
  for (int b = 0; b < nbins; ++b)
    if (!(k_edges[b + 1] > k_edges[b])) return fail(RF_ERR_ARG, "rf_r2c_power: bin edges must be strictly increasing");
  if (c->world > 1) return fail(RF_ERR_STATE, "rf_r2c_power: single-GPU contexts only");
  if (!c->field_is_real) return fail(RF_ERR_STATE, "rf_r2c_power: no real-space field");
  if (!(c->spacing > 0)) return fail(RF_ERR_STATE, "rf_r2c_power: call rf_set_power first (grid spacing)");
  return c->dtype == RF_F32 ? r2c_power_impl<float>(c, nbins, k_edges, out_pk, out_counts)
                            : r2c_power_impl<double>(c, nbins, k_edges, out_pk, out_counts);
}

// ------------------------------------------------------------------------------------------
// transform.Plan.execute on host arrays
// ------------------------------------------------------------------------------------------
template <typename T>
static int plan_impl(int nx, int ny, int nz, int inverse, int packed, const void* hin, void* hout) {
  const long long nzh = nz / 2 + 1;
  const long long ncplx = packed ? (long long)nx * ny * nzh : (long long)nx * ny * nz;
  const long long nreal = (long long)nx * ny * nz;
  const size_t cbytes = sizeof(cpx<T>) * ncplx, rbytes = sizeof(T) * nreal;
  void *a = nullptr, *b = nullptr;
  RF_CUDA(cudaMalloc(&a, cbytes));
  cudaStream_t s = nullptr;
  const double N3 = (double)nx * ny * nz;
  int e = RF_OK;
  auto cleanup = [&]() { if (a) cudaFree(a); if (b) cudaFree(b); };
#define PL_CUDA(x) do { cudaError_t e__ = (cudaError_t)(x); if (e__ != cudaSuccess) { cleanup(); return fail(RF_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e__)); } } while (0)
#define PL_RUN(x) do { e = (x); if (e) { cleanup(); return e; } } while (0)
  if (!packed) {
    PL_CUDA(cudaMemcpy(a, hin, cbytes, cudaMemcpyHostToDevice));
    const long long plane = (long long)ny * nz;
    PL_RUN(run_generic_axis<T>(nullptr, a, a, 0, inverse, nx, ny, nz, nz, 1, nz, 1, plane, plane, (T)1, s));
    PL_RUN(run_generic_axis<T>(nullptr, a, a, 0, inverse, ny, nx, nz, plane, 1, plane, 1, nz, nz, (T)1, s));
    PL_RUN(run_generic_axis<T>(nullptr, a, a, 0, inverse, nz, (long long)nx * ny, 1, nz, 0, nz, 0, 1, 1,
                               inverse ? (T)(1.0 / N3) : (T)1, s));
    PL_CUDA(cudaMemcpy(hout, a, cbytes, cudaMemcpyDeviceToHost));
  } else if (inverse) {
    PL_CUDA(cudaMalloc(&b, rbytes));
    PL_CUDA(cudaMemcpy(a, hin, cbytes, cudaMemcpyHostToDevice));
    const long long plane = (long long)ny * nzh;
    PL_RUN(run_generic_axis<T>(nullptr, a, a, 0, 1, nx, ny, nzh, nzh, 1, nzh, 1, plane, plane, (T)1, s));
    PL_RUN(run_generic_axis<T>(nullptr, a, a, 0, 1, ny, nx, nzh, plane, 1, plane, 1, nzh, nzh, (T)1, s));
    PL_RUN(run_generic_axis<T>(nullptr, a, b, 1, 1, nz, (long long)nx * ny, 1, nzh, 0, nz, 0, 1, 1,
                               (T)(1.0 / N3), s));
    PL_CUDA(cudaMemcpy(hout, b, rbytes, cudaMemcpyDeviceToHost));
  } else {
    PL_CUDA(cudaMalloc(&b, rbytes));
    PL_CUDA(cudaMemcpy(b, hin, rbytes, cudaMemcpyHostToDevice));
    const long long plane = (long long)ny * nzh;
    PL_RUN(run_generic_axis<T>(nullptr, b, a, 2, 0, nz, (long long)nx * ny, 1, nz, 0, nzh, 0, 1, 1, (T)1, s));
    PL_RUN(run_generic_axis<T>(nullptr, a, a, 0, 0, ny, nx, nzh, plane, 1, plane, 1, nzh, nzh, (T)1, s));
    PL_RUN(run_generic_axis<T>(nullptr, a, a, 0, 0, nx, ny, nzh, nzh, 1, nzh, 1, plane, plane, (T)1, s));
    PL_CUDA(cudaMemcpy(hout, a, cbytes, cudaMemcpyDeviceToHost));
  }
  cleanup();
  return RF_OK;
#undef PL_CUDA
#undef PL_RUN
}

extern "C" int rf_plan_execute(int nx, int ny, int nz, int dtype, int inverse, int packed,
                               const void* host_in, void* host_out) {
  if (!host_in || !host_out) return fail(RF_ERR_ARG, "rf_plan_execute: null argument");
  if (nx <= 0 || ny <= 0 || nz <= 0 || (nx % 2) || (ny % 2) || (nz % 2))
    return fail(RF_ERR_ARG, "All shape dimensions must be even.");
  return dtype == RF_F32 ? plan_impl<float>(nx, ny, nz, inverse, packed, host_in, host_out)
                         : plan_impl<double>(nx, ny, nz, inverse, packed, host_in, host_out);
}

// ------------------------------------------------------------------------------------------
// stage functions on host arrays
// ------------------------------------------------------------------------------------------
template <class F>
static int with_device_copy(void* host, size_t bytes, F&& body) {
  void* d = nullptr;
  RF_CUDA(cudaMalloc(&d, bytes));
  cudaError_t e = cudaMemcpy(d, host, bytes, cudaMemcpyHostToDevice);
  int r = RF_OK;
  if (e == cudaSuccess) r = body(d);
  if (e == cudaSuccess && r == RF_OK) e = cudaMemcpy(host, d, bytes, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return fail(RF_ERR_CUDA, cudaGetErrorString(e));
  return r;
}
static int grid_for(long long n) { return (int)std::min<long long>((n + 255) / 256, 148LL * 16); }

extern "C" int rf_stage_fill_log10k(int nx, int ny, int nz, int packed, int dtype, double spacing,
                                    void* host) {
  if (!host) return fail(RF_ERR_ARG, "null argument");
  const int ncols = packed ? nz / 2 + 1 : nz;
  const long long total = (long long)nx * ny * ncols;
  const double lambda0 = spacing / 6.283185307179586476925286766559;
  const size_t es = dtype == RF_F32 ? 8 : 16;
  return with_device_copy(host, es * total, [&](void* d) {
    if (dtype == RF_F32)
      stage_fill_log10k_kernel<float><<<grid_for(total), 256>>>((cpx<float>*)d, nx, ny, nz, ncols, lambda0);
    else
      stage_fill_log10k_kernel<double><<<grid_for(total), 256>>>((cpx<double>*)d, nx, ny, nz, ncols, lambda0);
    RF_CUDA(cudaGetLastError());
    return (int)RF_OK;
  });
}

extern "C" int rf_stage_tabulate(long long nelem, int dtype, int n, const double* log10k,
                                 const double* sigma, void* host) {
  if (!host || !log10k || !sigma || n < 2) return fail(RF_ERR_ARG, "bad argument");
  double *dx = nullptr, *df = nullptr;
  RF_CUDA(cudaMalloc((void**)&dx, sizeof(double) * n));
  RF_CUDA(cudaMalloc((void**)&df, sizeof(double) * n));
  cudaMemcpy(dx, log10k, sizeof(double) * n, cudaMemcpyHostToDevice);
  cudaMemcpy(df, sigma, sizeof(double) * n, cudaMemcpyHostToDevice);
  const size_t es = dtype == RF_F32 ? 8 : 16;
  int r = with_device_copy(host, es * nelem, [&](void* d) {
    if (dtype == RF_F32)
      stage_tabulate_kernel<float><<<grid_for(nelem), 256>>>((cpx<float>*)d, nelem, n, dx, df);
    else
      stage_tabulate_kernel<double><<<grid_for(nelem), 256>>>((cpx<double>*)d, nelem, n, dx, df);
    RF_CUDA(cudaGetLastError());
    return (int)RF_OK;
  });
  cudaFree(dx);
  cudaFree(df);
  return r;
}

extern "C" int rf_stage_randomize(long long nelem, int dtype, uint64_t seed, void* host) {
  if (!host) return fail(RF_ERR_ARG, "null argument");
  const size_t es = dtype == RF_F32 ? 8 : 16;
  return with_device_copy(host, es * nelem, [&](void* d) {
    if (dtype == RF_F32)
      stage_randomize_kernel<float><<<grid_for(nelem), 256>>>((cpx<float>*)d, nelem, (uint32_t)seed, (uint32_t)(seed >> 32));
    else
      stage_randomize_kernel<double><<<grid_for(nelem), 256>>>((cpx<double>*)d, nelem, (uint32_t)seed, (uint32_t)(seed >> 32));
    RF_CUDA(cudaGetLastError());
    return (int)RF_OK;
  });
}

extern "C" int rf_stage_symmetrize(int nx, int ny, int nz, int packed, int dtype, void* host) {
  if (!host) return fail(RF_ERR_ARG, "null argument");
  if ((nx % 2) || (ny % 2) || (nz % 2)) return fail(RF_ERR_ARG, "All shape dimensions must be even.");
  const int ncols = packed ? nz / 2 + 1 : nz;
  const long long total = (long long)nx * ny * ncols;
  const size_t es = dtype == RF_F32 ? 8 : 16;
  void* src = nullptr;
  RF_CUDA(cudaMalloc(&src, es * total));
  cudaMemcpy(src, host, es * total, cudaMemcpyHostToDevice);
  int r = with_device_copy(host, es * total, [&](void* d) {
    if (dtype == RF_F32)
      stage_symmetrize_kernel<float><<<grid_for(total), 256>>>((const cpx<float>*)src, (cpx<float>*)d, nx, ny, nz, packed);
    else
      stage_symmetrize_kernel<double><<<grid_for(total), 256>>>((const cpx<double>*)src, (cpx<double>*)d, nx, ny, nz, packed);
    RF_CUDA(cudaGetLastError());
    return (int)RF_OK;
  });
  cudaFree(src);
  return r;
}

template <typename T>
static int stage_density_impl(long long nelem, int nz, int lognormal, double sigma,
                              const double* growth, const double* rho, void* d) {
  if (lognormal && !(sigma > 0)) {
    double *parts = nullptr, *out = nullptr, h[2];
    RF_CUDA(cudaMalloc((void**)&parts, sizeof(double) * 2 * 1024));
    RF_CUDA(cudaMalloc((void**)&out, sizeof(double) * 2));
    stats_partial_kernel<T><<<1024, 256>>>((const T*)d, nelem, parts);
    stats_final_kernel<<<1, 256>>>(parts, 1024, out);
    cudaError_t e = cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    cudaFree(parts); cudaFree(out);
    RF_CUDA(e);
    double mean = h[0] / (double)nelem;
    sigma = std::sqrt(std::max(0.0, h[1] / (double)nelem - mean * mean));
  }
  std::vector<T> a(nz), b(nz);
  for (int z = 0; z < nz; ++z) {
    if (lognormal) {
      double t = 1.0 + (sigma * growth[z]) * (sigma * growth[z]);
      a[z] = (T)(std::sqrt(std::log(t)) / sigma);
      b[z] = (T)(rho[z] / std::sqrt(t));
    } else {
      a[z] = (T)growth[z];
      b[z] = (T)rho[z];
    }
  }
  T *da = nullptr, *db = nullptr;
  RF_CUDA(cudaMalloc((void**)&da, sizeof(T) * nz));
  RF_CUDA(cudaMalloc((void**)&db, sizeof(T) * nz));
  cudaMemcpy(da, a.data(), sizeof(T) * nz, cudaMemcpyHostToDevice);
  cudaMemcpy(db, b.data(), sizeof(T) * nz, cudaMemcpyHostToDevice);
  density_kernel<T><<<grid_for(nelem), 256>>>((T*)d, nelem, nz, da, db, lognormal);
  cudaError_t e = cudaDeviceSynchronize();
  cudaFree(da); cudaFree(db);
  RF_CUDA(e);
  return RF_OK;
}

extern "C" int rf_stage_density(long long nelem, int nz, int dtype, int lognormal, double sigma,
                                const double* growth, const double* rho, void* host) {
  if (!host || !growth || !rho || nz <= 0 || nelem <= 0 || nelem % nz)
    return fail(RF_ERR_ARG, "rf_stage_density: bad argument");
  const size_t es = dtype == RF_F32 ? 4 : 8;
  return with_device_copy(host, es * nelem, [&](void* d) {
    return dtype == RF_F32 ? stage_density_impl<float>(nelem, nz, lognormal, sigma, growth, rho, d)
                           : stage_density_impl<double>(nelem, nz, lognormal, sigma, growth, rho, d);
  });
}

// ------------------------------------------------------------------------------------------
// rf_density_noalloc: the density epilogue on caller-owned DEVICE arrays without any allocation
// (for callers that cannot call cudaMalloc: stream capture, serving loops).  No counterpart in
// the reference.  Same arithmetic as
// rf_stage_density / stage_density_impl above; the per-z tables and the partial sums of the
// sigma measurement live in ONE file-scope device array that every call on a device shares.
// ------------------------------------------------------------------------------------------
struct NoAllocScratch {
  double a[RF_NOALLOC_MAX_NZ];      // in: growth[z]    -> out: table a[z] (as double)
  double b[RF_NOALLOC_MAX_NZ];      // in: rho_mean[z]  -> out: table b[z]
  double parts[2 * 1024];           // per-CTA (sum, sumsq) of the sigma measurement
  double sums[2];                   // their fixed-order total
};
// Host code must never use this name as a pointer: it is a host-side stand-in for the symbol.
// noalloc_scratch() below asks the runtime for the device address, once per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

static const int kNoAllocMaxDevices = 64;
static NoAllocScratch* g_noalloc_ptr[kNoAllocMaxDevices];   // device address on each device, 0 = not asked yet

static int noalloc_scratch(NoAllocScratch** out) {
  int dev = -1;
  RF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(RF_ERR_ARG, "rf_density_noalloc: device ordinal out of range");
  NoAllocScratch* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    void* sym = nullptr;
    RF_CUDA(cudaGetSymbolAddress(&sym, g_noalloc_scratch));   // with `dev` current: this device's copy
    if (!sym) return fail(RF_ERR_CUDA, "rf_density_noalloc: cudaGetSymbolAddress returned a null address");
    p = (NoAllocScratch*)sym;
    __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);   // same value from every thread
  }
  *out = p;
  return RF_OK;
}

// one CTA of 256 threads: the total of the partial sums in stats_final_kernel's order (when sigma
// is to be measured), then growth / rho_mean -> the tables a, b of density_kernel, in place.
// Every scratch slot read here was written earlier in the same call: nothing has to start at zero.
template <typename T>
__global__ void noalloc_tables_kernel(NoAllocScratch* s, int nz, long long nelem, int lognormal, double sigma) {
  __shared__ double r1[256], r2[256];
  __shared__ double sig;
  if (lognormal && !(sigma > 0)) {
    double s1 = 0, s2 = 0;
    for (int i = threadIdx.x; i < 1024; i += 256) {
      s1 += s->parts[2 * i];
      s2 += s->parts[2 * i + 1];
    }
    r1[threadIdx.x] = s1;
    r2[threadIdx.x] = s2;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) {
        r1[threadIdx.x] += r1[threadIdx.x + o];
        r2[threadIdx.x] += r2[threadIdx.x + o];
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      s->sums[0] = r1[0];
      s->sums[1] = r2[0];
      const double mean = r1[0] / (double)nelem;
      sig = sqrt(fmax(0.0, r2[0] / (double)nelem - mean * mean));
    }
  } else if (threadIdx.x == 0) {
    sig = sigma;
  }
  __syncthreads();
  for (int z = threadIdx.x; z < nz; z += 256) {
    const double g = s->a[z], rm = s->b[z];
    if (lognormal) {
      const double t = 1.0 + (sig * g) * (sig * g);
      s->a[z] = (double)(T)(sqrt(log(t)) / sig);
      s->b[z] = (double)(T)(rm / sqrt(t));
    } else {
      s->a[z] = (double)(T)g;
      s->b[z] = (double)(T)rm;
    }
  }
}

// density_kernel with separate input / output arrays (they may be the same) and double tables
template <typename T>
__global__ void noalloc_density_kernel(const T* in, T* out, long long n, int nz,
                                       const NoAllocScratch* __restrict__ s, int lognormal) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const int z = (int)(i % nz);
    const T d = in[i];
    const T az = (T)s->a[z], bz = (T)s->b[z];
    out[i] = lognormal ? exp(d * az) * bz : ((T)1 + az * d) * bz;
  }
}

template <typename T>
static int density_noalloc_impl(NoAllocScratch* s, const void* in, void* out, long long nelem, int nz,
                                int lognormal, double sigma, cudaStream_t st) {
  if (lognormal && !(sigma > 0)) {
    stats_partial_kernel<T><<<1024, 256, 0, st>>>((const T*)in, nelem, s->parts);   // writes all 1024 slots
    RF_CUDA(cudaGetLastError());
  }
  noalloc_tables_kernel<T><<<1, 256, 0, st>>>(s, nz, nelem, lognormal, sigma);
  RF_CUDA(cudaGetLastError());
  noalloc_density_kernel<T><<<grid_for(nelem), 256, 0, st>>>((const T*)in, (T*)out, nelem, nz, s, lognormal);
  RF_CUDA(cudaGetLastError());
  return RF_OK;
}

extern "C" int rf_density_noalloc(const void* dev_delta, void* dev_out, long long nelem, int nz, int dtype,
                                  int lognormal, double sigma, const double* growth, const double* rho,
                                  void* cuda_stream) {
  if (!dev_delta || !dev_out || !growth || !rho || nz <= 0 || nelem <= 0 || nelem % nz)
    return fail(RF_ERR_ARG, "rf_density_noalloc: bad argument");
  if (dtype != RF_F32 && dtype != RF_F64) return fail(RF_ERR_ARG, "rf_density_noalloc: dtype must be RF_F32 or RF_F64");
  if (nz > RF_NOALLOC_MAX_NZ)
    return fail(RF_ERR_ARG, "rf_density_noalloc: nz = " + std::to_string(nz) + " exceeds RF_NOALLOC_MAX_NZ = " +
                                std::to_string(RF_NOALLOC_MAX_NZ) + " (the static scratch holds two tables of that length)");
  NoAllocScratch* s = nullptr;
  int e = noalloc_scratch(&s);
  if (e) return e;
  cudaStream_t st = (cudaStream_t)cuda_stream;
  // the tables start as growth / rho_mean; the copies are ordered on `st` before the kernels
  RF_CUDA(cudaMemcpyAsync(s->a, growth, sizeof(double) * nz, cudaMemcpyHostToDevice, st));
  RF_CUDA(cudaMemcpyAsync(s->b, rho, sizeof(double) * nz, cudaMemcpyHostToDevice, st));
  return dtype == RF_F32 ? density_noalloc_impl<float>(s, dev_delta, dev_out, nelem, nz, lognormal, sigma, st)
                         : density_noalloc_impl<double>(s, dev_delta, dev_out, nelem, nz, lognormal, sigma, st);
}
