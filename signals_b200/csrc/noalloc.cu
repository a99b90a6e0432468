// noalloc.cu -- one entry point for callers that are not allowed to allocate device memory.
//
// Some users of this library cannot call cudaMalloc at the point where they need a decode: services that hand
// the whole device to a single arena at start-up and treat any later driver allocation as a bug, latency-critical
// loops that cannot afford the implicit device-wide synchronisation of cudaMalloc / cudaFree, and code that has
// made one warm-up call and from then on must stay free of allocations.  Every other entry point of the library
// grows internal buffers on demand; this one never does.  It still makes ordinary, non-allocating CUDA runtime
// calls (cudaGetDevice, cudaFuncSetAttribute, a kernel launch; cudaGetSymbolAddress on the first call), so it must
// NOT be called from inside a cudaLaunchHostFunc callback, where every CUDA API call is forbidden.
// It is the chained (textbook-backtrace) variant of the stand-alone viterbi -- the one kernel whose backpointer
// scratch is otherwise allocated by the library -- with that scratch held in a fixed array that is part of the
// loaded module:
//
//   * the array is SGB_NOALLOC_SCRATCH_BYTES (4 MiB) and every GPU that loads the library carries it until the
//     process exits, whether or not this entry point is ever used;
//   * a problem needs T * S bytes of it (one uint8 backpointer per bin and state); larger problems are refused
//     with an error, never silently truncated;
//   * inputs and outputs are device pointers supplied by the caller, on the current device;
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device
//     current) and remembers it.  Make that first call before any stream capture begins on the device;
//   * all calls on a device share the one array and it keeps whatever the previous call left in it.  Nothing in
//     it has to start at zero: the kernel writes every backpointer of bins 1..T-1 before the backtrace reads
//     them and never reads bin 0's.  Two calls whose kernels overlap in time on the same GPU (different
//     streams, or different host threads) race on the array and corrupt each other's result: serialise them.
#include "../../include/signals_b200.h"
#include "kernels.h"
#include <mutex>
#include <string>

#define SGB_NOALLOC_SCRATCH_BYTES (4u << 20)

// File-scope device array.  In host code this NAME is only a stand-in symbol, not a device address: it must
// never be passed to a kernel directly -- the address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char sgb_noalloc_scratch[SGB_NOALLOC_SCRATCH_BYTES];

namespace {
std::mutex g_setup_mu;
unsigned char* g_scratch_addr[64] = {nullptr};       // per device, filled by the first call on that device

int noalloc_fail(const std::string& msg)
{
    sgb_internal_set_error(msg.c_str());
    return 1;
}
}  // namespace

extern "C" SGB_API int sgb_viterbi_chained_noalloc(const double* d_emission, int64_t nrow_e,
                                                   const double* d_transition, const int32_t* d_observations,
                                                   int64_t T, int32_t S, int32_t init_mode, int32_t* d_path_out,
                                                   void* stream)
{
    if (T < 0 || nrow_e < 0) return noalloc_fail("negative size");
    if (S < 1 || S > 128) return noalloc_fail("viterbi: number of states must be in 1..128");
    if (T == 0) return 0;
    if (nrow_e < 1) return noalloc_fail("viterbi: the emission matrix has no rows, so no observation index can be valid");
    if (!d_emission || !d_transition || !d_observations || !d_path_out) return noalloc_fail("NULL device pointer");
    // compared without forming T * S, which would wrap for absurd T and slip past the limit
    if ((unsigned long long)T > SGB_NOALLOC_SCRATCH_BYTES / (unsigned)S)
        return noalloc_fail("sgb_viterbi_chained_noalloc: T*S = " + std::to_string(T) + " x " + std::to_string(S) +
                            (T <= (1ll << 40) ? " = " + std::to_string((unsigned long long)T * (unsigned)S) : std::string()) +
                            " bytes of backpointers exceed the fixed " + std::to_string(SGB_NOALLOC_SCRATCH_BYTES) +
                            "-byte scratch; use sgb_viterbi (which allocates) for a problem of this size");
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail(std::string("no usable CUDA device: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return noalloc_fail("device index out of range");
    unsigned char* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_setup_mu);
        if (!g_scratch_addr[dev]) {
            // once per device, with that device current: resolve the symbol to its address on THIS device
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, sgb_noalloc_scratch);
            if (e != cudaSuccess || !p)
                return noalloc_fail(std::string("cudaGetSymbolAddress(sgb_noalloc_scratch) failed: ") + cudaGetErrorString(e));
            g_scratch_addr[dev] = (unsigned char*)p;
        }
        scratch = g_scratch_addr[dev];
    }
    e = sgbk_viterbi(d_emission, nrow_e, d_transition, d_observations, T, S, init_mode, SGB_DECODE_CHAINED, scratch,
                     d_path_out, (cudaStream_t)stream);
    if (e != cudaSuccess) return noalloc_fail(std::string("viterbi launch failed: ") + cudaGetErrorString(e));
    return 0;
}
