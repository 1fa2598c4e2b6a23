// User-defined EvaluateModelOnPoint functors: the README's solver concept lets any
// class supply `double EvaluateModelOnPoint(const Model&, int i) const` (README.md:105-154 of the
// reference); the three example solvers have hand-written sweeps (score_kernels.cu, sweep_tc_kernel.cu),
// everything else opts in by handing over the body of that function as CUDA source,
//
//     __device__ double rlb_eval(const double* model, const double* point);   // squared residual
//
// which is compiled here with NVRTC for sm_100a together with a K x N MSAC sweep and a per-point error
// kernel, and loaded through the driver API. libnvrtc and libcuda are opened at run time (dlopen), so the
// library still loads - and its CPU-side tests still run - on a machine without a driver.
//
// The sweep follows ScoreModel / ComputeScore (ransac.h:295-309: score = sum (thr < e) ? thr : e, NaN
// propagating, f64) with the same [splits][K] partials as the built-in sweeps, so finalize_kernel (validity,
// NaN rule, lowest-id argmin) is shared; the inlier pass evaluates the errors with the user's functor and
// then runs the ordered compaction of inlier_kernel on them.
#include <dlfcn.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "rlb_internal.cuh"

namespace rlb {

namespace {

// ---- minimal NVRTC / driver API surface (resolved with dlsym) -----------------------------------
typedef struct _nvrtcProgram* nvrtcProgram;
typedef int nvrtcResult;
typedef int CUresult;
typedef struct CUmod_st* CUmodule;
typedef struct CUfunc_st* CUfunction;
typedef struct CUstream_st* CUstream;

struct Nvrtc {
  void* lib = nullptr;
  nvrtcResult (*CreateProgram)(nvrtcProgram*, const char*, const char*, int, const char* const*, const char* const*);
  nvrtcResult (*CompileProgram)(nvrtcProgram, int, const char* const*);
  nvrtcResult (*GetProgramLogSize)(nvrtcProgram, size_t*);
  nvrtcResult (*GetProgramLog)(nvrtcProgram, char*);
  nvrtcResult (*GetCUBINSize)(nvrtcProgram, size_t*);
  nvrtcResult (*GetCUBIN)(nvrtcProgram, char*);
  nvrtcResult (*DestroyProgram)(nvrtcProgram*);
  const char* (*GetErrorString)(nvrtcResult);
};
struct Driver {
  void* lib = nullptr;
  CUresult (*ModuleLoadData)(CUmodule*, const void*);
  CUresult (*ModuleUnload)(CUmodule);
  CUresult (*ModuleGetFunction)(CUfunction*, CUmodule, const char*);
  CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream,
                           void**, void**);
  CUresult (*GetErrorString)(CUresult, const char**);
};

template <class Fn>
bool sym(void* lib, const char* name, Fn* out) {
  *out = reinterpret_cast<Fn>(dlsym(lib, name));
  return *out != nullptr;
}

const Nvrtc* nvrtc(std::string* err) {
  static Nvrtc n;
  static std::once_flag once;
  static std::string load_err;
  std::call_once(once, [] {
    const char* names[] = {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12"};
    for (const char* nm : names)
      if ((n.lib = dlopen(nm, RTLD_NOW | RTLD_LOCAL)) != nullptr) break;
    if (!n.lib) {
      load_err = "libnvrtc.so.12 not found (NVRTC is needed for user-defined functors)";
      return;
    }
    const bool ok = sym(n.lib, "nvrtcCreateProgram", &n.CreateProgram) &&
                    sym(n.lib, "nvrtcCompileProgram", &n.CompileProgram) &&
                    sym(n.lib, "nvrtcGetProgramLogSize", &n.GetProgramLogSize) &&
                    sym(n.lib, "nvrtcGetProgramLog", &n.GetProgramLog) &&
                    sym(n.lib, "nvrtcGetCUBINSize", &n.GetCUBINSize) && sym(n.lib, "nvrtcGetCUBIN", &n.GetCUBIN) &&
                    sym(n.lib, "nvrtcDestroyProgram", &n.DestroyProgram) &&
                    sym(n.lib, "nvrtcGetErrorString", &n.GetErrorString);
    if (!ok) {
      load_err = "libnvrtc is missing an entry point";
      n.lib = nullptr;
    }
  });
  if (!n.lib) {
    *err = load_err;
    return nullptr;
  }
  return &n;
}

const Driver* driver(std::string* err) {
  static Driver d;
  static std::once_flag once;
  static std::string load_err;
  std::call_once(once, [] {
    d.lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!d.lib) {
      load_err = "libcuda.so.1 not found (no NVIDIA driver on this machine)";
      return;
    }
    const bool ok = sym(d.lib, "cuModuleLoadData", &d.ModuleLoadData) && sym(d.lib, "cuModuleUnload", &d.ModuleUnload) &&
                    sym(d.lib, "cuModuleGetFunction", &d.ModuleGetFunction) &&
                    sym(d.lib, "cuLaunchKernel", &d.LaunchKernel) && sym(d.lib, "cuGetErrorString", &d.GetErrorString);
    if (!ok) {
      load_err = "libcuda is missing an entry point";
      d.lib = nullptr;
    }
  });
  if (!d.lib) {
    *err = load_err;
    return nullptr;
  }
  return &d;
}

std::string cu_error(const Driver* d, CUresult r) {
  const char* s = nullptr;
  d->GetErrorString(r, &s);
  char buf[96];
  snprintf(buf, sizeof(buf), "CUDA driver error %d (%s)", r, s ? s : "?");
  return buf;
}

// The kernels compiled around the user's functor. RLB_DIMS / RLB_MD are #defined in front of it.
const char* kKernelSource = R"SRC(
// RLB_HYBRID: the functor is the HybridSolver concept's EvaluateModelOnPoint(model, t, i)
// (README.md:164-208 of the reference; examples/hybrid_line_estimator.cc:100-109), i.e. it is told the data type.
#ifdef RLB_HYBRID
#define RLB_EVAL(m, t, p) rlb_eval_t(m, t, p)
#else
#define RLB_EVAL(m, t, p) rlb_eval(m, p)
#endif
// score = sum over live points of w_t * ((thr_t < e) ? thr_t : e) (ransac.h:306-309, hybrid_ransac.h:416-418),
// inlier count = #(e < thr_t) (ransac.h:319). One thread = one hypothesis; a CTA stages 128 points at a time in
// shared memory and every lane reads the same point (broadcast). blockIdx.y = split of the PADDED point axis:
// type 0 lives at [0, n0), type 1 at [off1, off1 + n1), off1 a multiple of 1024, so a chunk of 128 has one type.
extern "C" __global__ void __launch_bounds__(128)
rlb_custom_sweep(const double* __restrict__ soa, int npad, int n0, int off1, int n1,
                 const double* __restrict__ models, int k, double thr0, double thr1, double w0, double w1,
                 double* __restrict__ part_score, int* __restrict__ part_count, int pts_per_split, int want_counts) {
  __shared__ double tile[RLB_DIMS][128];
  const int tid = threadIdx.x;
  const int kk = blockIdx.x * 128 + tid;
  const int split = blockIdx.y;
  double m[RLB_MD];
  const int src = kk < k ? kk : k - 1;
#pragma unroll
  for (int i = 0; i < RLB_MD; ++i) m[i] = models[(size_t)src * RLB_MD + i];
  const int p0 = split * pts_per_split;
  const int p1 = min(npad, p0 + pts_per_split);
  double s = 0.0;
  int c = 0;
  for (int base = p0; base < p1; base += 128) {
    const int t = (n1 > 0 && base >= off1) ? 1 : 0;
    const int cnt = max(0, min(128, (t ? off1 + n1 : n0) - base));
    if (cnt == 0) continue;  // padding between the types / behind the last one (uniform for the CTA)
    const double thr = t ? thr1 : thr0;
    const int g = base + tid;
#pragma unroll
    for (int d = 0; d < RLB_DIMS; ++d) tile[d][tid] = tid < cnt ? soa[(size_t)d * npad + g] : 0.0;
    __syncthreads();
    double sc = 0.0;
    for (int j = 0; j < cnt; ++j) {
      double p[RLB_DIMS];
#pragma unroll
      for (int d = 0; d < RLB_DIMS; ++d) p[d] = tile[d][j];
      const double e = RLB_EVAL(m, t, p);
      sc += (thr < e) ? thr : e;
      c += (e < thr) ? 1 : 0;
    }
    s += sc * (t ? w1 : w0);
    __syncthreads();
  }
  if (kk < k) {
    part_score[(size_t)split * k + kk] = s;
    if (want_counts) part_count[(size_t)split * k + kk] = c;
  }
}

// err[i] = functor(model, type of i, point i) for the live points; the padding is never read as live by the
// compaction
extern "C" __global__ void __launch_bounds__(256)
rlb_custom_errors(const double* __restrict__ soa, int npad, int n0, int off1, int n1,
                  const double* __restrict__ model, double* __restrict__ err) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= npad) return;
  const int t = (n1 > 0 && i >= off1) ? 1 : 0;
  const bool live = t ? (i - off1 < n1) : (i < n0);
  double e = 0.0;
  if (live) {
    double m[RLB_MD], p[RLB_DIMS];
#pragma unroll
    for (int j = 0; j < RLB_MD; ++j) m[j] = model[j];
#pragma unroll
    for (int d = 0; d < RLB_DIMS; ++d) p[d] = soa[(size_t)d * npad + i];
    e = RLB_EVAL(m, t, p);
  }
  err[i] = e;
}
)SRC";

}  // namespace

// One data type: n rows of `width` doubles -> channels [0, dims) of the segment [off, off + span) of the
// [dims][npad] channel-major array; channels >= width and rows >= n are zero.
__global__ void custom_repack_kernel(const double* __restrict__ raw, int n, int width, int dims, int npad, int off,
                                     int span, double* __restrict__ soa) {
  const int total = span * dims;
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x) {
    const int c = g / span, i = g - c * span;
    soa[static_cast<size_t>(c) * npad + off + i] = (i < n && c < width) ? raw[static_cast<size_t>(i) * width + c] : 0.0;
  }
}

cudaError_t launch_custom_repack(const double* raw, int n, int width, int dims, int npad, int off, int span,
                                 double* soa, cudaStream_t st, int* launches) {
  ++*launches;
  const int blocks = max(1, min(2048, (span * dims + 255) / 256));
  custom_repack_kernel<<<blocks, 256, 0, st>>>(raw, n, width, dims, npad, off, span, soa);
  return cudaGetLastError();
}

struct CustomModule {
  CUmodule mod = nullptr;
  CUfunction sweep = nullptr;
  CUfunction errors = nullptr;
  int dims = 0, md = 0;
};

// Compiles `src` (+ the kernels above) to an sm_100a cubin. Needs no device. On failure *err holds the
// NVRTC log, so a typo in the functor reads like a compiler error.
bool custom_compile(const char* src, int dims, int md, bool hybrid, std::string* cubin, std::string* err) {
  const Nvrtc* n = nvrtc(err);
  if (!n) return false;
  // In front of the user's source: the sizes, and rlb_rcp(x) - the branch-free f64 reciprocal of the built-in sweep
  // (seed + two Newton steps, <= 1 ulp, x == 0 gives the signed infinity of the seed). `a / b` in f64 compiles to a
  // call with a slow path; a functor that may deviate from IEEE division by an ulp writes a * rlb_rcp(b) instead and
  // sweeps about twice as fast. Opt-in: nothing uses it unless the functor does.
  char defs[900];
  snprintf(defs, sizeof(defs),
           "#define RLB_DIMS %d\n#define RLB_MD %d\n%s"
           "__device__ __forceinline__ double rlb_rcp(double x) {\n"
           "  double r0;\n"
           "  asm(\"rcp.approx.ftz.f64 %%0, %%1;\" : \"=d\"(r0) : \"d\"(x));\n"
           "  double r = fma(fma(-x, r0, 1.0), r0, r0);\n"
           "  r = fma(fma(-x, r, 1.0), r, r);\n"
           "  return (x == 0.0) ? r0 : r;\n"
           "}\n",
           dims, md, hybrid ? "#define RLB_HYBRID 1\n" : "");
  // One compilation per distinct functor and process: a solver that is constructed per image / per query with the
  // same source pays NVRTC (10 - 15 ms) once. Bounded: the cache is dropped when it holds 64 entries.
  static std::mutex cache_mu;
  static std::unordered_map<std::string, std::string> cache;
  const std::string key = std::string(defs) + src;
  {
    std::lock_guard<std::mutex> lock(cache_mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *cubin = it->second;
      return true;
    }
  }
  std::string full = key + "\n" + kKernelSource;
  nvrtcProgram prog = nullptr;
  nvrtcResult r = n->CreateProgram(&prog, full.c_str(), "rlb_custom_functor.cu", 0, nullptr, nullptr);
  if (r != 0) {
    *err = std::string("nvrtcCreateProgram: ") + n->GetErrorString(r);
    return false;
  }
  // -fmad=false: a functor written like the reference's C++ evaluates like a strict-IEEE CPU build
  const char* opts[] = {"--gpu-architecture=sm_100a", "--std=c++17", "-fmad=false", "-lineinfo"};
  r = n->CompileProgram(prog, 4, opts);
  if (r != 0) {
    size_t len = 0;
    n->GetProgramLogSize(prog, &len);
    std::string log(len, '\0');
    if (len) n->GetProgramLog(prog, &log[0]);
    *err = std::string("NVRTC: ") + n->GetErrorString(r) + "\n" + log;
    n->DestroyProgram(&prog);
    return false;
  }
  size_t bytes = 0;
  n->GetCUBINSize(prog, &bytes);
  cubin->assign(bytes, '\0');
  if (bytes) n->GetCUBIN(prog, &(*cubin)[0]);
  n->DestroyProgram(&prog);
  if (!bytes) {
    *err = "NVRTC produced no cubin";
    return false;
  }
  {
    std::lock_guard<std::mutex> lock(cache_mu);
    if (cache.size() >= 64) cache.clear();
    cache[key] = *cubin;
  }
  return true;
}

CustomModule* custom_load(const std::string& cubin, int dims, int md, std::string* err) {
  const Driver* d = driver(err);
  if (!d) return nullptr;
  CustomModule* m = new CustomModule();
  m->dims = dims;
  m->md = md;
  CUresult r = d->ModuleLoadData(&m->mod, cubin.data());
  if (r == 0) r = d->ModuleGetFunction(&m->sweep, m->mod, "rlb_custom_sweep");
  if (r == 0) r = d->ModuleGetFunction(&m->errors, m->mod, "rlb_custom_errors");
  if (r != 0) {
    *err = cu_error(d, r);
    if (m->mod) d->ModuleUnload(m->mod);
    delete m;
    return nullptr;
  }
  return m;
}

void custom_unload(CustomModule* m) {
  if (!m) return;
  std::string err;
  const Driver* d = driver(&err);
  if (d && m->mod) d->ModuleUnload(m->mod);
  delete m;
}

bool launch_custom_sweep(const CustomModule* m, const DeviceData& d, const SweepArgs& a, int pts_per_split,
                         cudaStream_t st, int* launches, std::string* err) {
  const Driver* drv = driver(err);
  if (!drv) return false;
  ++*launches;
  const double* soa = d.soa64;
  int npad = d.npad, n0 = d.n[0], off1 = d.off[1], n1 = (d.num_types > 1) ? d.n[1] : 0, k = a.k, pps = pts_per_split,
      wc = a.want_counts;
  const double* models = a.models;
  double thr0 = a.thr[0], thr1 = a.thr[1], w0 = a.w[0], w1 = a.w[1];
  double* ps = a.part_score;
  int* pc = a.part_count;
  void* args[] = {&soa, &npad, &n0, &off1, &n1, &models, &k, &thr0, &thr1, &w0, &w1, &ps, &pc, &pps, &wc};
  const CUresult r = drv->LaunchKernel(m->sweep, (a.k + 127) / 128, a.n_splits, 1, 128, 1, 1, 0,
                                       reinterpret_cast<CUstream>(st), args, nullptr);
  if (r != 0) {
    *err = cu_error(drv, r);
    return false;
  }
  return true;
}

bool launch_custom_errors(const CustomModule* m, const DeviceData& d, const double* d_model, double* d_err,
                          cudaStream_t st, int* launches, std::string* err) {
  const Driver* drv = driver(err);
  if (!drv) return false;
  ++*launches;
  const double* soa = d.soa64;
  int npad = d.npad, n0 = d.n[0], off1 = d.off[1], n1 = (d.num_types > 1) ? d.n[1] : 0;
  void* args[] = {&soa, &npad, &n0, &off1, &n1, &d_model, &d_err};
  const CUresult r = drv->LaunchKernel(m->errors, (d.npad + 255) / 256, 1, 1, 256, 1, 1, 0,
                                       reinterpret_cast<CUstream>(st), args, nullptr);
  if (r != 0) {
    *err = cu_error(drv, r);
    return false;
  }
  return true;
}

}  // namespace rlb
