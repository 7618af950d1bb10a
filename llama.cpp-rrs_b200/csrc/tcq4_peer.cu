// tcq4_peer.cu -- cross-GPU barrier for the N-sharded MUL_MAT exchange (no counterpart in the fork: its TCQ4 path is
// single-GPU, ggml-cuda.cu:2195-2200).
//
// Consumers of an N-sharded C must not start before every rank's column block has landed in their copy.  One tiny kernel per
// barrier: every rank announces its next epoch in each peer's flag array (release, system scope: the stores of everything
// before it in stream order -- epilogue peer stores, multicast stores, copy-engine pushes the stream has waited for --
// are visible first) and spins until every peer has announced the same epoch to it (acquire).  The epoch lives on the
// device and advances by one per call, so the launch is identical every time: capturable, replayable.
// ~3-5 us per barrier against ~25 us for the symmetric-memory barrier of the framework (36 per Qwen3-4B step).
#include "tcq4_kernels.h"

namespace tcq4 {

struct PeerFlags {
    int *p[MAX_PEERS];      // flags of every rank (peer-mapped): int32[MAX_PEERS + 1], zero before the first barrier
};

__global__ void __launch_bounds__(32) tcq4_peer_barrier_kernel(const PeerFlags f, int n, int rank) {
    const int lane = threadIdx.x;
    int *mine = f.p[rank];
    int epoch = 0;
    if (lane == 0) {
        epoch = mine[MAX_PEERS] + 1;          // slot MAX_PEERS: this rank's own count of barriers
        mine[MAX_PEERS] = epoch;
    }
    epoch = __shfl_sync(0xffffffffu, epoch, 0);
    __threadfence_system();
    if (lane < n) {
        asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(f.p[lane] + rank), "r"(epoch) : "memory");
        int seen;
        do {
            asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(seen) : "l"(mine + lane) : "memory");
        } while (seen - epoch < 0);
    }
    __syncwarp();
    __threadfence_system();
}

cudaError_t launch_peer_barrier(int *const *flags, int n, int rank, cudaStream_t stream) {
    if (!flags || n < 1 || n > MAX_PEERS || rank < 0 || rank >= n) return cudaErrorInvalidValue;
    PeerFlags f;
    for (int i = 0; i < MAX_PEERS; i++) f.p[i] = flags[i < n ? i : 0];
    tcq4_peer_barrier_kernel<<<1, 32, 0, stream>>>(f, n, rank);
    return cudaGetLastError();
}

}  // namespace tcq4
