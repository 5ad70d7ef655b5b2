// noalloc_api.cu -- rt_intensity_noalloc: the one entry point of librt_b200.so that allocates NOTHING.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  It is rt_intensity
// (== intensity_calc, iteration_lvg.cpp:163-185 / :503-526; rt_api.cu) with the two scratch buffers that rt_intensity
// cudaMallocs per call -- the device copy of the layer's populations and the {result, dust cross-sections} block --
// replaced by ONE file-scope device array of fixed size.  Every GPU that loads the library holds that array
// (RT_NOALLOC_SCRATCH_BYTES, 8.3 KB) until the process exits.
//
// Rules this file follows:
//  * `g_noalloc_scratch` is a host-side stand-in in host code, not a device address: the device address is fetched once per
//    device, with that device current, by cudaGetSymbolAddress (noalloc_scratch_ptr) and handed to the kernel;
//  * the array is shared by every call on a device and keeps the last call's contents: each call overwrites everything the
//    kernel reads (populations [0, N), the result slot cleared to 0, the cross-sections) before the launch;
//  * calls that overlap on one GPU (two host threads, two rt_model handles on the same device) RACE on the array: serialise them;
//  * no cudaMalloc / cudaFree / allocating library routine anywhere on the path; N beyond RT_NOALLOC_MAX_LEVELS is refused.
#include "rt_internal.h"
#include <cmath>
#include <cstdio>
#include <mutex>

#define NOALLOC_OUT_SLOTS 16 // {J, sigma[RT_MAX_DUST]} padded
static_assert(1 + RT_MAX_DUST <= NOALLOC_OUT_SLOTS, "result block too small");
static_assert(RT_NOALLOC_SCRATCH_BYTES == (RT_NOALLOC_MAX_LEVELS + NOALLOC_OUT_SLOTS) * sizeof(double), "header and array disagree");

static __device__ __align__(256) double g_noalloc_scratch[RT_NOALLOC_MAX_LEVELS + NOALLOC_OUT_SLOTS];

#define NOALLOC_MAX_DEVICES 64
static double* g_scratch_ptr[NOALLOC_MAX_DEVICES]; // device address of g_noalloc_scratch on each GPU, nullptr = not fetched yet
static std::mutex g_scratch_mutex;

// set-up, once per device, with that device current
static cudaError_t noalloc_scratch_ptr(int device, double** out)
{
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice; // the caller makes the device current first
        double* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

// dust_component::absorption(energy), dust_model.cpp:473-490, from the host copies the model keeps
static double noalloc_dust_sigma(const rt_model* m, int c, double energy)
{
    const std::vector<double>&en = m->h_dust_en[c], &ab = m->h_dust_abs[c];
    const int n = (int)en.size();
    if (energy < en[0]) return ab[0] * pow(en[0] / energy, m->h_dust_exp[c]);
    if (energy > en[n - 1]) return ab[n - 1];
    int l = 0, r = n - 1;
    while (r - l > 1) {
        const int i = l + ((r - l) >> 1);
        if (en[i] < energy) l = i; else r = i;
    }
    const double deriv = (ab[l + 1] - ab[l]) / (en[l + 1] - en[l]); // dust_model.cpp:516-524
    return ab[l] + deriv * (energy - en[l]);
}

#define NA_CK(call)                                                                                              \
    do {                                                                                                         \
        cudaError_t e_ = (call);                                                                                 \
        if (e_ != cudaSuccess) {                                                                                 \
            char buf_[512];                                                                                      \
            snprintf(buf_, sizeof buf_, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));    \
            m->last_error = buf_;                                                                                \
            return RT_ERR_CUDA;                                                                                  \
        }                                                                                                        \
    } while (0)

extern "C" rt_status rt_intensity_noalloc(rt_model* m, int lay, int up, int low, const double* pop, double* intensity)
{
    if (!m) return RT_ERR_INVALID;
    m->last_error.clear();
    if (!m->has_cloud) { m->last_error = "no cloud uploaded"; return RT_ERR_NO_CLOUD; }
    const DevModel& d = m->d;
    const int N = d.N;
    if (N > RT_NOALLOC_MAX_LEVELS) {
        char buf[160];
        snprintf(buf, sizeof buf, "rt_intensity_noalloc: nb_lev = %d exceeds RT_NOALLOC_MAX_LEVELS = %d (fixed scratch array); use rt_intensity", N, RT_NOALLOC_MAX_LEVELS);
        m->last_error = buf;
        return RT_ERR_UNSUPPORTED;
    }
    if (lay < 0 || lay >= d.L || up < 0 || up >= N || low < 0 || low >= N || !pop || !intensity) return RT_ERR_INVALID;
    NA_CK(cudaSetDevice(m->device));
    double* scratch = nullptr;
    NA_CK(noalloc_scratch_ptr(m->device, &scratch));
    // which line(s): as rt_intensity (iteration_lvg.cpp:503-526: look the line up in line_list; not found -> 0)
    int which = 0, u1 = up, l1 = low, u2 = up, l2 = low;
    if (d.scheme == RT_SCHEME_LINE_OVERLAP) {
        bool found = false;
        for (size_t i = 0; i < m->h_grp.size() / 5 && !found; i++) {
            const int* g = &m->h_grp[5 * i];
            for (int k = 0; k < g[0]; k++)
                if (g[2 + 2 * k] == low && g[1 + 2 * k] == up) {
                    found = true;
                    if (g[0] == 2) { which = k + 1; u1 = g[1]; l1 = g[2]; u2 = g[3]; l2 = g[4]; }
                    break;
                }
        }
        if (!found) { *intensity = 0.; return RT_OK; }
    }
    double* pop_dev = scratch;                          // [0, N): the layer's populations
    double* out_dev = scratch + RT_NOALLOC_MAX_LEVELS;  // {J, sigma[ncomp]}
    double block[NOALLOC_OUT_SLOTS] = {0.};             // block[0] = 0: the result slot starts cleared
    const double e1 = m->h_energy[u1] - m->h_energy[l1];
    for (int c = 0; c < d.ncomp; c++) block[1 + c] = noalloc_dust_sigma(m, c, e1);
    NA_CK(cudaMemcpyAsync(pop_dev, pop, sizeof(double) * N, cudaMemcpyHostToDevice, m->stream));
    NA_CK(cudaMemcpyAsync(out_dev, block, sizeof block, cudaMemcpyHostToDevice, m->stream));
    const std::vector<double>& A = m->h_einst;
    launch_intensity_single(d, m->stream, lay, u1, l1, A[(size_t)u1 * N + l1], A[(size_t)l1 * N + u1], u2, l2,
                            A[(size_t)u2 * N + l2], A[(size_t)l2 * N + u2], which, pop_dev, out_dev);
    NA_CK(cudaGetLastError());
    NA_CK(cudaMemcpyAsync(intensity, out_dev, sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    NA_CK(cudaStreamSynchronize(m->stream));
    return RT_OK;
}
