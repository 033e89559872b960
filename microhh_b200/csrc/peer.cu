// Peer-memory data path of the y-slab decomposition: one process per GPU, every rank's exchange arena is
// mapped into its neighbours' address spaces (CUDA IPC over NVLink / NVSwitch), and the kernels of this file move data with
// plain stores to peer addresses and synchronise with flags in the same arenas — no pack / unpack passes, no host waits.
//
//  * halo_push_kernel: the north/south part of Boundary_cyclic::exec on MPI (src/boundary_cyclic.cxx:139-174): my top
//    interior rows go straight into the north neighbour's south ghost rows, my bottom interior rows into the south
//    neighbour's north ghost rows, for up to 8 fields per launch.
//  * peer_barrier_kernel: all ranks have finished (and made visible) everything stream-ordered before it; placed between the
//    FFT pass that writes the all-to-all chunks into the peers' receive buffers (Transpose::exec_xy / exec_yx,
//    src/transpose.cxx:163-217, fused into the transforms) and the pass that reads them.
//
// Flag protocol (64-bit epochs, monotonically increasing, identical on all ranks because all ranks issue the same sequence):
//   ready : "every kernel I launched before this exchange is complete" -> my neighbours may overwrite my ghost rows
//   data  : "my rows for this epoch are in your ghost rows"
// A rank writes its flags INTO THE NEIGHBOUR'S arena (slots from_south / from_north) and spins only on its own memory.
// Spins give up after ~20 s (clock64) and raise the arena's error word, so a lost peer cannot hang the GPU.
#include "sys.h"
#include "ctx.h"

namespace mhh
{
namespace
{
    constexpr int SLOT_READY_FROM_SOUTH = 0, SLOT_READY_FROM_NORTH = 1, SLOT_DATA_FROM_SOUTH = 2, SLOT_DATA_FROM_NORTH = 3;
    constexpr int SLOT_BARRIER = 8;      // + rank (< 8)
    constexpr int SLOT_ERROR = 16;
    constexpr int SLOT_COUNTER = 24;     // local block counter of the push kernel
    constexpr long long SPIN_LIMIT = 40000000000LL;     // ~20 s of SM clocks

    __device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
    { asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory"); }
    __device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
    {
        unsigned long long v;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
        return v;
    }
    // waits until *p >= epoch; false on timeout (the error word is raised)
    __device__ __forceinline__ bool spin_until(const unsigned long long* p, const unsigned long long epoch, unsigned long long* err)
    {
        const long long t0 = clock64();
        while (ld_acquire_sys(p) < epoch)
        {
            if (clock64() - t0 > SPIN_LIMIT) { atomicExch(err, 1ULL); return false; }
            __nanosleep(64);
        }
        return true;
    }

    struct HaloJob
    {
        double* fld[8];
        int nf, nk;          // fields, levels per field (kcells, or 1 for 2-D planes)
    };

    // grid: (ceil(icells / 128), jgc, nk * nf)
    __global__ void __launch_bounds__(128)
    halo_push_kernel(const HaloJob job, const long north_delta, const long south_delta,   // byte offsets: peer arena - my arena
                     unsigned long long* __restrict__ flags, const unsigned long long epoch, const GridD gd)
    {
        unsigned long long* nflags = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(flags) + north_delta);
        unsigned long long* sflags = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(flags) + south_delta);
        const bool first = (blockIdx.x | blockIdx.y | blockIdx.z) == 0;
        __shared__ int ok;
        if (threadIdx.x == 0)
        {
            // everything stream-ordered before this kernel is complete on this rank: the neighbours may write my ghost rows
            if (first)
            {
                st_release_sys(nflags + SLOT_READY_FROM_SOUTH, epoch);
                st_release_sys(sflags + SLOT_READY_FROM_NORTH, epoch);
            }
            ok = spin_until(flags + SLOT_READY_FROM_NORTH, epoch, flags + SLOT_ERROR)
              && spin_until(flags + SLOT_READY_FROM_SOUTH, epoch, flags + SLOT_ERROR);
        }
        __syncthreads();
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        const int g = blockIdx.y;
        const int f = blockIdx.z / job.nk, k = blockIdx.z - f * job.nk;
        if (ok && i < gd.icells)
        {
            double* fld = job.fld[f];
            const long base = i + (long)k * gd.ijcells;
            double* nfld = reinterpret_cast<double*>(reinterpret_cast<char*>(fld) + north_delta);
            double* sfld = reinterpret_cast<double*>(reinterpret_cast<char*>(fld) + south_delta);
            nfld[base + (long)g * gd.icells] = fld[base + (long)(gd.jend - gd.jgc + g) * gd.icells];             // -> its south ghost rows
            sfld[base + (long)(gd.jend + g) * gd.icells] = fld[base + (long)(gd.jstart + g) * gd.icells];      // -> its north ghost rows
        }
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0)
        {
            unsigned* counter = reinterpret_cast<unsigned*>(flags + SLOT_COUNTER);
            const unsigned total = gridDim.x * gridDim.y * gridDim.z;
            if (atomicAdd(counter, 1u) == total - 1)
            {
                // last block: all rows of this rank are on their way; tell both neighbours, then wait for theirs
                *counter = 0u;
                __threadfence_system();
                st_release_sys(nflags + SLOT_DATA_FROM_SOUTH, epoch);
                st_release_sys(sflags + SLOT_DATA_FROM_NORTH, epoch);
                spin_until(flags + SLOT_DATA_FROM_NORTH, epoch, flags + SLOT_ERROR);
                spin_until(flags + SLOT_DATA_FROM_SOUTH, epoch, flags + SLOT_ERROR);
            }
        }
    }

    struct PeerDeltas { long d[8]; };

    // one CTA; thread r < world signals rank r and waits for rank r's signal
    __global__ void peer_barrier_kernel(const PeerDeltas deltas, const int world, const int me,
                                        unsigned long long* __restrict__ flags, const unsigned long long epoch)
    {
        __threadfence_system();
        const int r = threadIdx.x;
        if (r < world)
        {
            unsigned long long* pf = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(flags) + deltas.d[r]);
            st_release_sys(pf + SLOT_BARRIER + me, epoch);
            spin_until(flags + SLOT_BARRIER + r, epoch, flags + SLOT_ERROR);
        }
    }
}  // namespace

int peer_halo_push(Ctx* c, int nf, double* const* flds, int is2d)
{
    const GridD& gd = c->gd;
    if (!c->peer.flags) { c->last_error = "peer arena not registered (mhh_peer_register)"; return MHH_ERR_STATE; }
    if (nf < 1 || nf > 8) { c->last_error = "mhh_peer_halo_push: 1..8 fields per call"; return MHH_ERR_ARG; }
    HaloJob job;
    job.nf = nf; job.nk = is2d ? 1 : gd.kcells;
    const char* lo = reinterpret_cast<const char*>(c->peer.base);
    for (int q = 0; q < nf; ++q)
    {
        const char* p = reinterpret_cast<const char*>(flds[q]);
        if (p < lo || p + sizeof(double) * (is2d ? gd.ijcells : gd.ncells) > lo + c->peer.bytes)
        { c->last_error = "mhh_peer_halo_push: field is not inside the registered peer arena"; return MHH_ERR_ARG; }
        job.fld[q] = flds[q];
    }
    const unsigned long long epoch = ++c->peer.halo_epoch;
    const dim3 b(128), g((gd.icells + 127) / 128, gd.jgc, job.nk * nf);
    MHH_LAUNCH(c, "halo_push_kernel", halo_push_kernel<<<g, b, 0, c->stream>>>(
        job, c->peer.delta[c->peer.north], c->peer.delta[c->peer.south], c->peer.flags, epoch, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int peer_barrier(Ctx* c)
{
    if (!c->peer.flags) { c->last_error = "peer arena not registered (mhh_peer_register)"; return MHH_ERR_STATE; }
    PeerDeltas d;
    for (int r = 0; r < 8; ++r) d.d[r] = r < c->peer.world ? c->peer.delta[r] : 0;
    const unsigned long long epoch = ++c->peer.barrier_epoch;
    MHH_LAUNCH(c, "peer_barrier_kernel", peer_barrier_kernel<<<1, 32, 0, c->stream>>>(d, c->peer.world, c->peer.me, c->peer.flags, epoch));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}

int peer_error(Ctx* c, int* err)
{
    if (!c->peer.flags) { *err = 0; return MHH_OK; }
    unsigned long long v = 0;
    MHH_CUDA_CHECK(cudaMemcpyAsync(&v, c->peer.flags + SLOT_ERROR, sizeof(v), cudaMemcpyDeviceToHost, c->stream));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    *err = v != 0;
    return MHH_OK;
}

}  // namespace mhh
