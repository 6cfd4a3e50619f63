// ns_noalloc.inl -- ONE entry point that allocates nothing (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// nsgpu_minmax_noalloc is the VecMin / VecMax x 4 part of the time step (FS.cpp:554-557; k_correct's NaN-propagating reduction) on two
// caller arrays.  Every context-based entry point keeps its reduction scratch (Reducer::partials, ::counter) and its fields in
// cudaMalloc'ed memory; here ALL scratch -- the staging copies of u and v, the per-block partial results, the block counter and the
// result -- lives in one file-scope static __device__ array of fixed size, which every GPU that loads libnsgpu.so holds until exit:
//   NSGPU_NOALLOC_MAX_CELLS = 2^20 cells  ->  2 x 8 MiB staging + 2 MiB partials + 64 B = 18.0 MiB per GPU.
// In host code the array's NAME is only a host-side stand-in: its device address is fetched once per device, with that device
// current, by cudaGetSymbolAddress and cached; the kernel receives that pointer.  The array is shared by every call and keeps the last
// call's contents: the block counter is cleared at the start of every call, everything else is overwritten before it is read.  Calls
// that overlap on one GPU (two host threads, or two processes' worth of streams) race on it -- callers serialise.
// No cudaMalloc, cudaFree, cudaMallocHost, stream or event creation: synchronous copies and one launch on the legacy default stream.

namespace nsgpu {

constexpr long long NOALLOC_MAX_CELLS = 1ll << 20;
constexpr size_t NOALLOC_OFF_V = (size_t)NOALLOC_MAX_CELLS;                     // doubles
constexpr size_t NOALLOC_OFF_PART = 2 * (size_t)NOALLOC_MAX_CELLS;              // 4 x RED_MAX_BLOCKS partial maxima
constexpr size_t NOALLOC_OFF_CNT = NOALLOC_OFF_PART + 4 * (size_t)RED_MAX_BLOCKS;   // block counter (one 8-byte slot)
constexpr size_t NOALLOC_OFF_OUT = NOALLOC_OFF_CNT + 4;                         // 4 results
constexpr size_t NOALLOC_DOUBLES = NOALLOC_OFF_OUT + 4;

static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_DOUBLES];

// out[0..3] = max(-u), max(u), max(-v), max(v) over n cells; a non-finite value anywhere gives +inf on all four (as k_correct)
__global__ void __launch_bounds__(256) k_minmax_scratch(const double* __restrict__ u, const double* __restrict__ v, long long n,
                                                        Reducer red, double* __restrict__ out) {
  double acc[4] = {OpMax::id(), OpMax::id(), OpMax::id(), OpMax::id()};
  const double inf = 1.0 / 0.0;
  for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    const double un = u[k], vn = v[k];
    const bool bad = !(isfinite(un) && isfinite(vn));
    acc[0] = fmax(acc[0], bad ? inf : -un); acc[1] = fmax(acc[1], bad ? inf : un);
    acc[2] = fmax(acc[2], bad ? inf : -vn); acc[3] = fmax(acc[3], bad ? inf : vn);
  }
  grid_reduce<4, OpMax>(acc, red, out);
}

// device address of the static array, once per device (index = CUDA device ordinal)
static double* noalloc_scratch_ptr(int device, std::string& err) {
  constexpr int MAXDEV = 64;
  static double* ptr[MAXDEV] = {};
  static std::mutex mu;
  if (device < 0 || device >= MAXDEV) { err = "nsgpu_minmax_noalloc: device ordinal outside [0, 64)"; return nullptr; }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[device]) {
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);   // the current device is `device` (set by the caller below)
    if (e != cudaSuccess || !p) { err = std::string("nsgpu_minmax_noalloc: cudaGetSymbolAddress -> ") + cudaGetErrorString(e); cudaGetLastError(); return nullptr; }
    ptr[device] = p;
  }
  return ptr[device];
}

}  // namespace nsgpu

extern "C" {

long long nsgpu_noalloc_max_cells(void) { return nsgpu::NOALLOC_MAX_CELLS; }

int nsgpu_minmax_noalloc(int32_t device, const double* u, const double* v, int64_t n, double* out4) {
  using namespace nsgpu;
  std::string& err = g_create_error;   // read back with nsgpu_last_error(NULL)
  if (!u || !v || !out4 || n < 1) { err = "nsgpu_minmax_noalloc: null array or n < 1"; return NSGPU_ERR_ARG; }
  if (n > NOALLOC_MAX_CELLS) {
    char buf[200];
    snprintf(buf, sizeof buf, "nsgpu_minmax_noalloc: n = %lld exceeds the static scratch (NSGPU_NOALLOC_MAX_CELLS = %lld cells)", (long long)n, NOALLOC_MAX_CELLS);
    err = buf;
    return NSGPU_ERR_ARG;
  }
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) { err = std::string("nsgpu_minmax_noalloc: cudaSetDevice -> ") + cudaGetErrorString(e); cudaGetLastError(); return NSGPU_ERR_CUDA; }
  double* s = noalloc_scratch_ptr(device, err);
  if (!s) return NSGPU_ERR_CUDA;
  // the counter must start at zero (a call that died between launch and completion would leave it elsewhere); staging, partials and
  // the result are written before they are read
  e = cudaMemset(s + NOALLOC_OFF_CNT, 0, 32);
  if (e == cudaSuccess) e = cudaMemcpy(s, u, (size_t)n * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(s + NOALLOC_OFF_V, v, (size_t)n * sizeof(double), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { err = std::string("nsgpu_minmax_noalloc: copy in -> ") + cudaGetErrorString(e); cudaGetLastError(); return NSGPU_ERR_CUDA; }
  Reducer red{s + NOALLOC_OFF_PART, reinterpret_cast<unsigned*>(s + NOALLOC_OFF_CNT), nullptr};
  const int blocks = (int)std::min<long long>((n + 255) / 256, 148 * 8);
  k_minmax_scratch<<<blocks, 256>>>(s, s + NOALLOC_OFF_V, (long long)n, red, s + NOALLOC_OFF_OUT);
  e = cudaGetLastError();
  double h[4];
  if (e == cudaSuccess) e = cudaMemcpy(h, s + NOALLOC_OFF_OUT, sizeof h, cudaMemcpyDeviceToHost);   // synchronises with the kernel
  if (e != cudaSuccess) { err = std::string("nsgpu_minmax_noalloc: kernel / copy out -> ") + cudaGetErrorString(e); cudaGetLastError(); return NSGPU_ERR_CUDA; }
  out4[0] = -h[0]; out4[1] = h[1]; out4[2] = -h[2]; out4[3] = h[3];   // umin, umax, vmin, vmax
  return 0;
}

}  // extern "C"
