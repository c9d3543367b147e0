// noalloc.cu -- mpic_kinetic_energy_static: the kinetic-energy / total-charge reduction of mpic_diagnostics for callers that
// must not allocate device memory.  (No counterpart in the reference, which has neither particles nor diagnostics.)
//
// Who cannot allocate, in practice, around this library:
//   * a host code that keeps its particle arrays in its OWN device arena (its own pool, a torch caching allocator, an MPI code
//     that reserved the whole HBM for a 1e9-particle slab at start-up).  It cannot spare the ~1 GB of planes, accumulators and
//     sort tables mpic_create would add just to ask "what is the kinetic energy now", and next to a full arena cudaMalloc
//     simply fails;
//   * a diagnostics hook that fires between the steps of a running peer ring: cudaMalloc / cudaFree synchronise the whole
//     device and serialise against every stream of the process, and the neighbours' kernels are spinning on this rank's flags
//     meanwhile, so an allocation in the hook stalls all ranks, not one.
// mpic_diagnostics itself is of no use to them: it needs a handle, and a handle owns cudaMalloc'ed state.
//
// What the reduction needs besides its inputs (the four particle arrays) and its outputs (two doubles) is a work area: one
// pair of partial sums per thread block, a ticket the blocks draw to find out who finishes last, and two result slots the last
// block fills.  Here that work area is the file-scope array g_kin_work below, MPIC_KINETIC_STATIC_SCRATCH_BYTES long, part of
// the library image: every GPU that loads libminipic_b200.so carries these 9.5 KB until the process exits, used or not.
//
// Host code must not hand `g_kin_work` to a kernel: on the host that name is only the handle the runtime looks the array up
// by, not a device address.  The address is asked for once per device, with that device current, through
// cudaGetSymbolAddress (kin_work_address below) and it is that pointer the launch gets.
#include "common.cuh"
#include <mutex>

namespace {

constexpr int KIN_BLOCKS = 592;                    // == launch_kinetic (particles.cu): same blocks, same partials, same sum order
constexpr int KIN_THREADS = 256;
constexpr int KIN_RESULT = 0;                      // doubles [0,2): kinetic sum, charge sum
constexpr int KIN_TICKET = 2;                      // double slot 2, first 4 bytes: ticket counter
constexpr int KIN_PART = 4;                        // doubles [4, 4 + 2*KIN_BLOCKS): block partials
constexpr int KIN_HEAD_BYTES = KIN_PART * 8;       // what has to be zero when a launch starts
constexpr int KIN_WORK_DOUBLES = MPIC_KINETIC_STATIC_SCRATCH_BYTES / 8;
static_assert(KIN_PART + 2 * KIN_BLOCKS <= KIN_WORK_DOUBLES, "work area too small for the block partials");
static_assert(MPIC_KINETIC_STATIC_SCRATCH_BYTES % 256 == 0, "keep the array a whole number of 256-byte lines");

static __device__ __align__(256) double g_kin_work[KIN_WORK_DOUBLES];

// Same arithmetic, thread-to-particle map and summation order as kinetic_kernel (particles.cu), on bare pointers instead of
// the handle's Particles view; the division by |q/m| that mpic_diagnostics does on the host is done by the last block.
__global__ void __launch_bounds__(KIN_THREADS) kinetic_static_kernel(const double* __restrict__ ux, const double* __restrict__ uy,
                                                                      const double* __restrict__ uz, const double* __restrict__ q,
                                                                      int64_t n, double abs_qm, double* __restrict__ work) {
    __shared__ double sh[KIN_THREADS], sq[KIN_THREADS];
    double s = 0.0, c = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double u2 = ux[i] * ux[i] + uy[i] * uy[i] + uz[i] * uz[i];
        s += fabs(q[i]) * (u2 / (sqrt(1.0 + u2) + 1.0));
        c += q[i];
    }
    sh[threadIdx.x] = s; sq[threadIdx.x] = c;
    __syncthreads();
    for (int w = KIN_THREADS / 2; w > 0; w >>= 1) {
        if (threadIdx.x < w) { sh[threadIdx.x] += sh[threadIdx.x + w]; sq[threadIdx.x] += sq[threadIdx.x + w]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        double* part = work + KIN_PART;
        unsigned int* ticket = reinterpret_cast<unsigned int*>(work + KIN_TICKET);
        part[2 * blockIdx.x] = sh[0]; part[2 * blockIdx.x + 1] = sq[0];
        __threadfence();
        if (atomicAdd(ticket, 1u) == gridDim.x - 1) {
            __threadfence();
            double t = 0.0, cs = 0.0;
            for (unsigned b = 0; b < gridDim.x; b++) { t += ((volatile double*)part)[2 * b]; cs += ((volatile double*)part)[2 * b + 1]; }
            work[KIN_RESULT] = t / abs_qm; work[KIN_RESULT + 1] = cs;
            *ticket = 0u;
        }
    }
}

// Device address of g_kin_work on the CURRENT device, looked up once per device and remembered.
constexpr int KIN_MAX_DEVICES = 64;
std::mutex g_kin_mu;
double* g_kin_addr[KIN_MAX_DEVICES] = { nullptr };

int kin_work_address(int dev, double** out) {
    std::lock_guard<std::mutex> lk(g_kin_mu);
    if (!g_kin_addr[dev]) {
        double* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_kin_work);   // device `dev` is current (checked by the caller)
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            mpic_set_error("mpic_kinetic_energy_static: cudaGetSymbolAddress on device %d failed: %s", dev, cudaGetErrorString(e));
            return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? MPIC_ENODEV : MPIC_ECUDA;
        }
        g_kin_addr[dev] = p;
    }
    *out = g_kin_addr[dev];
    return 0;
}

}  // namespace

#define NA_REQUIRE(cond, code, ...) do { if (!(cond)) { mpic_set_error(__VA_ARGS__); return code; } } while (0)

extern "C" int mpic_kinetic_energy_static(const double* ux, const double* uy, const double* uz, const double* q, int64_t n,
                                          double qm, double out[2], void* stream) {
    NA_REQUIRE(out, MPIC_EINVAL, "mpic_kinetic_energy_static: out is null");
    NA_REQUIRE(n >= 0, MPIC_EINVAL, "mpic_kinetic_energy_static: negative particle count");
    NA_REQUIRE(n <= MPIC_KINETIC_STATIC_MAX_PARTICLES, MPIC_EINVAL,
               "mpic_kinetic_energy_static: %lld particles exceed the maximum of %lld per call (the library's 32-bit slot range); "
               "reduce the population in pieces and add the results", (long long)n, (long long)MPIC_KINETIC_STATIC_MAX_PARTICLES);
    NA_REQUIRE(qm != 0.0 && qm == qm, MPIC_EINVAL, "mpic_kinetic_energy_static: charge/mass ratio must be a non-zero number");
    NA_REQUIRE(n == 0 || (ux && uy && uz && q), MPIC_EINVAL, "mpic_kinetic_energy_static: null particle array");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        mpic_set_error("no CUDA device: libminipic_b200 has no CPU fallback");
        return MPIC_ENODEV;
    }
    out[0] = 0.0; out[1] = 0.0;
    if (n == 0) return 0;
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    NA_REQUIRE(dev >= 0 && dev < KIN_MAX_DEVICES, MPIC_EINVAL, "mpic_kinetic_energy_static: device ordinal %d out of range", dev);
    // the arrays must be device memory of the device that is current: the work area used is that device's copy
    const double* arrs[4] = { ux, uy, uz, q };
    for (int k = 0; k < 4; k++) {
        cudaPointerAttributes at;
        CUDA_TRY(cudaPointerGetAttributes(&at, arrs[k]));
        const bool ok = at.type == cudaMemoryTypeManaged || (at.type == cudaMemoryTypeDevice && at.device == dev);
        NA_REQUIRE(ok, MPIC_EINVAL, "mpic_kinetic_energy_static: array %d is not device memory of the current device %d", k, dev);
    }
    double* work = nullptr;
    int r = kin_work_address(dev, &work);
    if (r) return r;
    cudaStream_t st = (cudaStream_t)stream;
    // The array is shared by all calls and still holds what the last one left.  Result slots and ticket must start at zero
    // (the kernel resets the ticket itself, but a launch that died half-way would leave it counting); the partials need no
    // clearing: every block overwrites its own pair before the last block reads any.
    CUDA_TRY(cudaMemsetAsync(work, 0, KIN_HEAD_BYTES, st));
    kinetic_static_kernel<<<KIN_BLOCKS, KIN_THREADS, 0, st>>>(ux, uy, uz, q, n, fabs(qm), work);
    {
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { mpic_set_error("mpic_kinetic_energy_static: launch failed: %s", cudaGetErrorString(e)); return MPIC_ECUDA; }
    }
    double h[2] = { 0.0, 0.0 };
    CUDA_TRY(cudaMemcpyAsync(h, work + KIN_RESULT, sizeof h, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    out[0] = h[0]; out[1] = h[1];
    return 0;
}
