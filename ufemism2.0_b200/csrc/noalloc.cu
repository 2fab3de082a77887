// ufe_b200_calc_pc_truncation_error_noalloc: the truncation-error kernel of geometry.cu (calc_pc_truncation_error,
// time_stepping/predictor_corrector_scheme.f90:327-361, 254-279) for callers that must not allocate device memory (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference: added at the request of the run's operators.
//
// Every byte of device memory the call needs -- the staged inputs, the per-vertex outputs and the three reduction cells -- lives in ONE file-scope
// __device__ array of fixed size (kNoAllocBytes, ~11 MiB, for at most UFE_B200_NOALLOC_MAX_VERTICES vertices).  Each GPU that loads the library holds
// that array until the process exits.  The entry point calls no cudaMalloc / cudaFree / allocating library routine: it neither builds the handle's
// geometry buffers (ensure_geometry) nor uses the timing events of UFE_LAUNCH.
//
// The array is shared by every call on a device and keeps the previous call's contents: the reduction cells are re-initialised by k_na_init on the
// stream before the kernel, all other regions are overwritten for [0, n) before they are read.  Two calls that overlap on the GPU (two handles / two
// host threads on one device) race on it -- serialise them.
#include <mutex>

#include "common.cuh"

namespace ufe {
namespace {

constexpr int kNaMax = UFE_B200_NOALLOC_MAX_VERTICES;
constexpr size_t kNaHeader = 256;                                   // [0] bits of the running maximum, [8] n_tot, [12] n_guilty
constexpr size_t kNaD = sizeof(double) * (size_t)kNaMax, kNaI = sizeof(int) * (size_t)kNaMax;   // both multiples of 256
constexpr size_t kNoAllocBytes = kNaHeader + 4 * kNaD + 3 * kNaI;  // 11 534 592 bytes
static_assert(kNaD % 256 == 0 && kNaI % 256 == 0, "regions of the scratch array stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

struct NaView {   // the regions of the scratch array, from its DEVICE address
  unsigned long long* etabits; int* counts;
  double *Hi_np1, *Hi_star, *frac, *tau;
  int *grounded, *gl_gr, *guilty;
  explicit NaView(unsigned char* base) {
    etabits = reinterpret_cast<unsigned long long*>(base); counts = reinterpret_cast<int*>(base + 8);
    unsigned char* p = base + kNaHeader;
    Hi_np1 = reinterpret_cast<double*>(p); p += kNaD; Hi_star = reinterpret_cast<double*>(p); p += kNaD;
    frac = reinterpret_cast<double*>(p); p += kNaD; tau = reinterpret_cast<double*>(p); p += kNaD;
    grounded = reinterpret_cast<int*>(p); p += kNaI; gl_gr = reinterpret_cast<int*>(p); p += kNaI; guilty = reinterpret_cast<int*>(p);
  }
};

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is asked for once per device, with that
// device current, and cached.
constexpr int kNaMaxDevices = 64;
std::mutex g_na_mutex;
unsigned char* g_na_ptr[kNaMaxDevices] = {nullptr};

int noalloc_scratch(int device, unsigned char** out) {
  UFE_CHECK(device >= 0 && device < kNaMaxDevices, "ufe_b200_calc_pc_truncation_error_noalloc: device %d outside [0, %d)", device, kNaMaxDevices);
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[device]) {
    void* p = nullptr;
    UFE_CUDA(cudaSetDevice(device));
    UFE_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    UFE_CHECK(p != nullptr, "ufe_b200_calc_pc_truncation_error_noalloc: cudaGetSymbolAddress returned a null address");
    g_na_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_na_ptr[device];
  return 0;
}

__global__ void k_na_init(unsigned long long* etabits, int* counts, unsigned long long init) {
  if (blockIdx.x == 0 && threadIdx.x == 0) { *etabits = init; counts[0] = 0; counts[1] = 0; }
}

// the arithmetic of k_pc_truncation_error (geometry.cu), statement for statement, on the regions of the scratch array
__global__ void k_na_truncation_error(int n, double zeta_t, double dt_n, double eps, const double* __restrict__ Hi_np1,
                                      const double* __restrict__ Hi_star, const int* __restrict__ grounded, const int* __restrict__ gl_gr,
                                      const double* __restrict__ fraction_gr, double* __restrict__ tau, int* __restrict__ guilty_rec,
                                      unsigned long long* etabits, int* counts) {
  double best = 0.0;
  int ntot = 0, ngui = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const double t = zeta_t * fabs(Hi_np1[i] - Hi_star[i]) / ((3.0 * zeta_t + 3.0) * dt_n);
    tau[i] = t;
    const bool full = fraction_gr[i] == 1.0;
    if (grounded[i] && !gl_gr[i] && full) best = fmax(best, t);
    if (full) {
      ++ntot;
      if (t > eps) { ++ngui; if (guilty_rec) guilty_rec[i] = guilty_rec[i] + 1; }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    best = fmax(best, __shfl_xor_sync(0xffffffffu, best, o));
    ntot += __shfl_xor_sync(0xffffffffu, ntot, o);
    ngui += __shfl_xor_sync(0xffffffffu, ngui, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (best > 0.0) atomicMax(etabits, (unsigned long long)__double_as_longlong(best));
    if (ntot) atomicAdd(&counts[0], ntot);
    if (ngui) atomicAdd(&counts[1], ngui);
  }
}

}  // namespace
}  // namespace ufe

using namespace ufe;

extern "C" int ufe_b200_calc_pc_truncation_error_noalloc(ufe_b200_handle h, int n, double zeta_t, double dt_n, double pc_eta_min, double pc_epsilon,
                                                         const double* Hi_np1, const double* Hi_star_np1, const int32_t* mask_grounded_ice,
                                                         const int32_t* mask_gl_gr, const double* fraction_gr, double* tau_np1,
                                                         int32_t* tau_n_guilty, double* eta_np1, int32_t* n_tot, int32_t* n_guilty) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && eta_np1 && n_tot && n_guilty, "ufe_b200_calc_pc_truncation_error_noalloc: NULL argument");
  UFE_CHECK(n >= 0, "ufe_b200_calc_pc_truncation_error_noalloc: n = %d is negative", n);
  UFE_CHECK(n <= kNaMax,
            "ufe_b200_calc_pc_truncation_error_noalloc: n = %d exceeds UFE_B200_NOALLOC_MAX_VERTICES = %d (the fixed scratch array); "
            "use ufe_b200_calc_pc_truncation_error", n, kNaMax);
  UFE_CHECK(dt_n > 0.0 && pc_eta_min > 0.0, "ufe_b200_calc_pc_truncation_error_noalloc: dt_n and pc_eta_min must be positive");
  if (n == 0) { *eta_np1 = pc_eta_min; *n_tot = 0; *n_guilty = 0; return 0; }
  UFE_CHECK(Hi_np1 && Hi_star_np1 && mask_grounded_ice && mask_gl_gr && fraction_gr, "ufe_b200_calc_pc_truncation_error_noalloc: NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  unsigned char* base = nullptr;
  UFE_TRY(noalloc_scratch(c->device, &base));
  UFE_CUDA(cudaSetDevice(c->device));
  NaView S(base);
  cudaStream_t st = c->stream;
  const size_t nd = sizeof(double) * (size_t)n, ni = sizeof(int) * (size_t)n;
  UFE_CUDA(cudaMemcpyAsync(S.Hi_np1, Hi_np1, nd, cudaMemcpyHostToDevice, st));
  UFE_CUDA(cudaMemcpyAsync(S.Hi_star, Hi_star_np1, nd, cudaMemcpyHostToDevice, st));
  UFE_CUDA(cudaMemcpyAsync(S.frac, fraction_gr, nd, cudaMemcpyHostToDevice, st));
  UFE_CUDA(cudaMemcpyAsync(S.grounded, mask_grounded_ice, ni, cudaMemcpyHostToDevice, st));
  UFE_CUDA(cudaMemcpyAsync(S.gl_gr, mask_gl_gr, ni, cudaMemcpyHostToDevice, st));
  if (tau_n_guilty) UFE_CUDA(cudaMemcpyAsync(S.guilty, tau_n_guilty, ni, cudaMemcpyHostToDevice, st));
  const unsigned long long init = (unsigned long long)__builtin_bit_cast(long long, pc_eta_min);
  k_na_init<<<1, 32, 0, st>>>(S.etabits, S.counts, init);          // the cells that must start from a known value (the array keeps the last call's)
  UFE_CUDA(cudaGetLastError());
  k_na_truncation_error<<<grid_for(c, n, 256), 256, 0, st>>>(n, zeta_t, dt_n, pc_epsilon, S.Hi_np1, S.Hi_star, S.grounded, S.gl_gr, S.frac, S.tau,
                                                             tau_n_guilty ? S.guilty : nullptr, S.etabits, S.counts);
  UFE_CUDA(cudaGetLastError());
  c->launches += 2;
  unsigned long long bits = 0; int counts[2] = {0, 0};
  if (tau_np1) UFE_CUDA(cudaMemcpyAsync(tau_np1, S.tau, nd, cudaMemcpyDeviceToHost, st));
  if (tau_n_guilty) UFE_CUDA(cudaMemcpyAsync(tau_n_guilty, S.guilty, ni, cudaMemcpyDeviceToHost, st));
  UFE_CUDA(cudaMemcpyAsync(&bits, S.etabits, sizeof(bits), cudaMemcpyDeviceToHost, st));
  UFE_CUDA(cudaMemcpyAsync(counts, S.counts, sizeof(counts), cudaMemcpyDeviceToHost, st));
  UFE_CUDA(cudaStreamSynchronize(st));
  *eta_np1 = __builtin_bit_cast(double, bits); *n_tot = (int32_t)counts[0]; *n_guilty = (int32_t)counts[1];
  return 0;
}
