// host_noalloc.cuh — nb_diagnostics_noalloc: Physics::compute_diagnostics without any allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). It is a variant of
// diag_exact_kernel (diag.cuh: the FP64 pair sum of Physics.hpp:148-154 and the per-body sums of :131-139) that
// reads the caller's DEVICE arrays in the reference's component layout and keeps its per-CTA partial rows in a
// file-scope device array instead of a buffer the context allocated. Part of nbody_b200.cu's translation unit
// (included there), so the array, the symbol lookup and the launches are in one .cu file.
//
// Scratch: g_noalloc_scratch, kNoAllocRows x kDiagValues doubles = 96 KiB, held by every GPU that loads the
// library until the process exits. Every call on a device shares it and nothing in it has to start at zero (each
// row the final pass reads was written by the same call's first kernel, ordered by the stream), but two calls
// that overlap on one GPU (different streams, or graphs replayed concurrently) race on it: serialise them.
#pragma once

// the first kernel writes one row of kDiagValues doubles per CTA of kDiagThreads bodies
constexpr long long kNoAllocMaxBodies = NB_NOALLOC_MAX_BODIES;
constexpr long long kNoAllocRows = (kNoAllocMaxBodies + nb::kDiagThreads - 1) / nb::kDiagThreads;
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocRows * nb::kDiagValues];

namespace nb {

// diag_exact_kernel over AoS inputs: pos/vel are n x dim doubles (dim 2: z = 0, as nb_upload stores it), same
// operations in the same order, partial rows to `part` (the device address of g_noalloc_scratch).
__global__ void __launch_bounds__(kDiagThreads) diag_exact_aos_kernel(const double* __restrict__ pos,
                                                                      const double* __restrict__ vel,
                                                                      const float* __restrict__ mass, long long n,
                                                                      int dim, double g, float softening,
                                                                      double* part) {
    __shared__ double sx[kDiagThreads], sy[kDiagThreads], sz[kDiagThreads], sm[kDiagThreads];
    __shared__ double red[kDiagThreads / 32][kDiagValues];
    const double eps2 = (double)softening * (double)softening;
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const bool valid = i < n;
    double xi[3] = {0, 0, 0}, vi[3] = {0, 0, 0};
    float mi = 0.f;
    if (valid) {
        for (int k = 0; k < dim; ++k) { xi[k] = pos[i * dim + k]; vi[k] = vel[i * dim + k]; }
        mi = mass[i];
    }
    double pe = 0.0;
    const long long j_start = blockIdx.x * (long long)blockDim.x;  // pairs i < j only
    for (long long j0 = j_start; j0 < n; j0 += kDiagThreads) {
        const long long j = j0 + threadIdx.x;
        if (j < n) {
            sx[threadIdx.x] = pos[j * dim];
            sy[threadIdx.x] = pos[j * dim + 1];
            sz[threadIdx.x] = dim == 3 ? pos[j * dim + 2] : 0.0;
            sm[threadIdx.x] = (double)mass[j];
        }
        __syncthreads();
        if (valid) {
            const int cnt = (int)min((long long)kDiagThreads, n - j0);
            for (int u = 0; u < cnt; ++u) {
                if (j0 + u <= i) continue;
                const double dx = sx[u] - xi[0], dy = sy[u] - xi[1], dz = sz[u] - xi[2];
                const double r = __dsqrt_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)), eps2));
                pe += __ddiv_rn(__dmul_rn(__dmul_rn(-g, (double)mi), sm[u]), r);
            }
        }
        __syncthreads();
    }
    DiagTerms dg;
#pragma unroll
    for (int k = 0; k < kDiagValues; ++k) dg.v[k] = 0.0;
    if (valid) {
        diag_accumulate(dg, xi, vi, mi, 0.0, g, false);
        dg.v[1] = pe;
    }
#pragma unroll
    for (int k = 0; k < 10; ++k) {
        double w = dg.v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][k] = w;
    }
    __syncthreads();
    if (threadIdx.x < 10) {
        double w = 0.0;
        for (int q = 0; q < kDiagThreads / 32; ++q) w += red[q][threadIdx.x];
        part[(long long)blockIdx.x * kDiagValues + threadIdx.x] = w;
    }
}

// sum_rows_kernel + the host arithmetic of finish_diag in one warp: rows added in ascending order, then the
// reference's Diagnostics (Physics.hpp:163-176) as nb_diag on the device.
__global__ void diag_finish_noalloc_kernel(const double* part, int rows, nb_diag* out) {
    __shared__ double t[kDiagValues];
    const int k = threadIdx.x;
    if (k < 10) {
        double acc = 0.0;
        for (int b = 0; b < rows; ++b) acc += part[(long long)b * kDiagValues + k];
        t[k] = acc;
    }
    __syncwarp();
    if (k != 0) return;
    const double M = t[8];
    const double vals[10] = {t[0], t[1], t[0] + t[1], t[2], t[3], t[4],
                             M > 0.0 ? t[5] / M : 0.0, M > 0.0 ? t[6] / M : 0.0, M > 0.0 ? t[7] / M : 0.0, M};
    bool ok = t[9] == 0.0;
    for (int q = 0; q < 10; ++q) ok = ok && finite_d(vals[q]);
    nb_diag d;
    d.kinetic = ok ? vals[0] : 0.0;
    d.potential = ok ? vals[1] : 0.0;
    d.energy = ok ? vals[2] : 0.0;
    for (int q = 0; q < 3; ++q) {
        d.momentum[q] = ok ? vals[3 + q] : 0.0;
        d.com[q] = ok ? vals[6 + q] : 0.0;
    }
    d.total_mass = ok ? M : 0.0;
    d.ok = ok ? 1 : 0;
    d._pad = 0;
    *out = d;
}

}  // namespace nb

namespace {

// Device address of g_noalloc_scratch per device. In host code the array's name is only a stand-in for the
// symbol: the address a kernel can use comes from cudaGetSymbolAddress, looked up once per device with that
// device current (each device holds its own copy of the array).
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mutex;
double* g_noalloc_ptr[kNoAllocMaxDevices] = {};

int noalloc_scratch(int device, double** out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        double* p = nullptr;
        NB_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) return fail(NB_ERR_CUDA, "nb_diagnostics_noalloc: cudaGetSymbolAddress returned a null address");
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return NB_OK;
}

struct DeviceScope {  // makes `device` current for the call and puts the caller's device back
    int prev = -1;
    cudaError_t enter(int device) {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) { prev = -1; return e; }
        return prev == device ? cudaSuccess : cudaSetDevice(device);
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

extern "C" int nb_diagnostics_noalloc(int device, int64_t n, int dim, const double* d_pos, const double* d_vel,
                                      const float* d_mass, double g, float softening, nb_diag* d_out,
                                      void* cuda_stream) {
    if (!d_out) return fail(NB_ERR_INVALID, "nb_diagnostics_noalloc: d_out is null");
    if (dim != 2 && dim != 3) return fail(NB_ERR_INVALID, "nb_diagnostics_noalloc: dim must be 2 or 3");
    if (n < 0) return fail(NB_ERR_INVALID, "nb_diagnostics_noalloc: negative body count");
    if (n > kNoAllocMaxBodies)
        return fail(NB_ERR_INVALID, "nb_diagnostics_noalloc: " + std::to_string((long long)n) +
                                        " bodies exceed NB_NOALLOC_MAX_BODIES (" + std::to_string(kNoAllocMaxBodies) +
                                        "), the size the static scratch array was built for; use nb_compute_diagnostics");
    if (n > 0 && (!d_pos || !d_vel || !d_mass))
        return fail(NB_ERR_INVALID, "nb_diagnostics_noalloc: null input array");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0)
        return fail(NB_ERR_NO_DEVICE, "nb_diagnostics_noalloc: no CUDA device");
    if (device < 0 || device >= count || device >= kNoAllocMaxDevices)
        return fail(NB_ERR_INVALID, "nb_diagnostics_noalloc: no such device");
    DeviceScope scope;
    NB_CUDA(scope.enter(device));
    double* part = nullptr;
    NB_TRY(noalloc_scratch(device, &part));
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    const long long rows = (n + nb::kDiagThreads - 1) / nb::kDiagThreads;  // <= kNoAllocRows by the check above
    if (rows > 0) {
        nb::diag_exact_aos_kernel<<<(unsigned int)rows, nb::kDiagThreads, 0, stream>>>(d_pos, d_vel, d_mass, n, dim, g,
                                                                                       softening, part);
        NB_CUDA(cudaGetLastError());
    }
    // n == 0: no rows -> all sums zero, ok = 1 (Physics.hpp:124-127)
    nb::diag_finish_noalloc_kernel<<<1, 32, 0, stream>>>(part, (int)rows, d_out);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}
