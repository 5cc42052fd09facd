// Instantiations and launcher of the row-resident (thread-block cluster + DSMEM + TMA) inverse kernel, acq_cluster.cuh.
#include "acq_cluster.cuh"
#include <cstdlib>

namespace b200
{
namespace cl
{
namespace
{
// B200_ACQ_CLUSTER: 0 = never use the row-resident kernel (two-kernel path), 1 = the variant that stages X_k by TMA bulk copies,
// 2 = operands in registers with the half-warp-per-column upper stage; unset / anything else = the default variant (operands in
// registers, grouped upper stage). Read once per process (tuning aid, like B200_NO_FIXED).
int env_variant()
{
    static const int v = [] {
        const char* e = std::getenv("B200_ACQ_CLUSTER");
        return e ? std::atoi(e) : -1;
    }();
    return v;
}

template <class G, int MODE>
cudaError_t launch_one(uint32_t n_rows, cudaStream_t st, const FastArgs& a)
{
    auto kernel = inv_cluster_wl_kernel<G, MODE>;
    static std::atomic<int> n_clusters_cached[kMaxDeviceSlots];  // per instantiation and device: co-resident clusters (0 = not yet)
    const int slot = device_slot();
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(G::THREADS);
    cfg.dynamicSmemBytes = G::smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = G::CS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n_clusters = slot >= 0 ? n_clusters_cached[slot].load(std::memory_order_relaxed) : 0;
    if (n_clusters <= 0)
        {
            cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::smem_bytes);
            if (e != cudaSuccess) return e;
            cfg.gridDim = dim3(G::CS * kNumSMs * G::CTAS_PER_SM);  // only the shape matters for the query
            e = cudaOccupancyMaxActiveClusters(&n_clusters, kernel, &cfg);
            if (e != cudaSuccess) return e;
            if (n_clusters <= 0) return cudaErrorLaunchOutOfResources;
            if (slot >= 0) n_clusters_cached[slot].store(n_clusters, std::memory_order_relaxed);
            if (std::getenv("B200_CL_VERBOSE")) fprintf(stderr, "inv_cluster CS=%d threads=%d smem=%zu: %d co-resident clusters\n", G::CS, G::THREADS, (size_t)G::smem_bytes, n_clusters);
        }
    static const uint32_t dbg = [] {
        const char* e = std::getenv("B200_CL_DBG");  // timing experiments only (results are wrong with any bit set)
        return e ? (uint32_t)std::atoi(e) : 0u;
    }();
    // a cluster looks the operand rows of at most kRowTab rows up at its start: larger searches go out in several launches
    const uint32_t per_launch = (uint32_t)n_clusters * (uint32_t)G::kRowTab;
    FastArgs part = a;
    for (uint32_t y0 = 0; y0 < n_rows; y0 += per_launch)
        {
            const uint32_t cnt = std::min(per_launch, n_rows - y0);
            part.y0 = a.y0 + y0;
            cfg.gridDim = dim3(std::min<uint32_t>((uint32_t)n_clusters, cnt) * G::CS);
            const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, part, cnt, dbg);
            if (e != cudaSuccess) return e;
        }
    return cudaSuccess;
}

template <class G>
cudaError_t launch_mode(int mode, uint32_t n_rows, cudaStream_t st, const FastArgs& a)
{
    if (mode == 1) return launch_one<G, 1>(n_rows, st, a);
    if (mode == 2) return launch_one<G, 2>(n_rows, st, a);
    return cudaErrorInvalidValue;
}

// N = 32 000 = (8 x 8) x (25 x 20) (BASELINE configs[1], Galileo E1 at 8 MS/s): a row is 256 kB = the shared memory of an SM pair.
// 8 x 8 lower blocks: 4 blocks = 32 pencils per warp in both lower passes; 16 warps x 2 columns of the 32 a CTA owns.
using G32k_g = GeoW<8, 8, 25, 20, 2, 512, 0, 1, 4>;  // operands in registers, grouped upper stage (default)
using G32k_r = GeoW<8, 8, 25, 20, 2, 512, 0>;        // operands in registers, one upper-stage column per half-warp (B200_ACQ_CLUSTER=2)
using G32k_w = GeoW<8, 8, 25, 20, 2, 512, 2>;        // X_k tiles by TMA bulk copies, two-stage ring (B200_ACQ_CLUSTER=1)
}  // namespace

bool cluster_preferred_plan(uint32_t N, uint32_t* ra_lo, uint32_t* rb_lo, uint32_t* ra_hi, uint32_t* rb_hi)
{
    if (env_variant() == 0) return false;
    if (N == 32000)
        {
            *ra_lo = 8, *rb_lo = 8, *ra_hi = 25, *rb_hi = 20;
            return true;
        }
    return false;
}

uint32_t cluster_tiles(uint32_t ra_lo, uint32_t rb_lo, uint32_t ra_hi, uint32_t rb_hi)
{
    if (env_variant() == 0) return 0;
    if (ra_lo == 8 && rb_lo == 8 && ra_hi == 25 && rb_hi == 20) return G32k_r::CS * G32k_r::NW;
    return 0;
}

cudaError_t launch_inv_cluster(uint32_t ra_lo, uint32_t rb_lo, uint32_t ra_hi, uint32_t rb_hi, int mode, uint32_t n_rows, cudaStream_t st, const FastArgs& a)
{
    if (ra_lo == 8 && rb_lo == 8 && ra_hi == 25 && rb_hi == 20)
        {
            if (env_variant() == 1) return launch_mode<G32k_w>(mode, n_rows, st, a);
            if (env_variant() == 2) return launch_mode<G32k_r>(mode, n_rows, st, a);
            return launch_mode<G32k_g>(mode, n_rows, st, a);
        }
    return cudaErrorInvalidValue;
}

}  // namespace cl
}  // namespace b200
