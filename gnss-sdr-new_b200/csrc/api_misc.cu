// Library-level entry points of libgnss_b200.so that are not tied to a handle.
#include "common.cuh"
#include <cmath>

namespace b200
{
char* last_error_buf()
{
    static thread_local char buf[512] = {0};
    return buf;
}

int select_device(int device)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(GNSS_B200_ERR_NO_DEVICE, "no CUDA device available (%s); libgnss_b200 has no CPU path", e == cudaSuccess ? "count = 0" : cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(GNSS_B200_ERR_INVALID, "device %d out of range [0,%d)", device, n);
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(GNSS_B200_ERR_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    return GNSS_B200_OK;
}

// Regularised lower incomplete gamma P(a, x): series for x < a+1, Lentz continued fraction otherwise.
static double gamma_p(double a, double x)
{
    if (x <= 0.0) return 0.0;
    const double pref = std::exp(-x + a * std::log(x) - std::lgamma(a));
    if (x < a + 1.0)
        {
            double term = 1.0 / a, sum = term;
            for (int k = 1; k < 200000; k++)
                {
                    term *= x / (a + k);
                    sum += term;
                    if (std::fabs(term) < std::fabs(sum) * 1e-17) break;
                }
            return sum * pref;
        }
    const double tiny = 1e-300;
    double b = x + 1.0 - a, c = 1.0 / tiny, d = 1.0 / b, f = d;
    for (int i = 1; i < 200000; i++)
        {
            const double an = -i * (i - a);
            b += 2.0;
            d = an * d + b;
            if (std::fabs(d) < tiny) d = tiny;
            c = b + an / c;
            if (std::fabs(c) < tiny) c = tiny;
            d = 1.0 / d;
            const double del = d * c;
            f *= del;
            if (std::fabs(del - 1.0) < 1e-16) break;
        }
    return 1.0 - pref * f;
}

// Inverse of P(a, .) at p by safeguarded Newton (stands in for boost::math::gamma_p_inv).
double gamma_p_inv(double a, double p)
{
    if (p <= 0.0) return 0.0;
    if (p >= 1.0) return INFINITY;
    double lo = 0.0, hi = a + 10.0;
    while (gamma_p(a, hi) < p) hi *= 2.0;
    double x = 0.5 * (lo + hi);
    const double lg = std::lgamma(a);
    for (int it = 0; it < 500; it++)
        {
            const double f = gamma_p(a, x) - p;
            (f > 0 ? hi : lo) = x;
            const double slope = std::exp((a - 1.0) * std::log(x) - x - lg);
            double xn = slope > 0 ? x - f / slope : 0.5 * (lo + hi);
            if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
            const bool done = std::fabs(xn - x) <= 1e-15 * std::fabs(x);
            x = xn;
            if (done) break;
        }
    return x;
}
}  // namespace b200

namespace
{
__global__ void fp32_peak_kernel(float* out, int iters)
{
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float m = 0.999f, c = 1e-3f;
    for (int i = 0; i < iters; i++)
        {
            a0 = fmaf(a0, m, c);
            a1 = fmaf(a1, m, c);
            a2 = fmaf(a2, m, c);
            a3 = fmaf(a3, m, c);
            a4 = fmaf(a4, m, c);
            a5 = fmaf(a5, m, c);
            a6 = fmaf(a6, m, c);
            a7 = fmaf(a7, m, c);
        }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
}  // namespace

extern "C"
{
    int gnss_b200_debug_fp32_peak(int32_t device, double* tflops)
    {
        B200_REQUIRE(tflops, GNSS_B200_ERR_INVALID, "fp32_peak: null");
        int rc = b200::select_device(device);
        if (rc) return rc;
        const int blocks = b200::kNumSMs * 8, threads = 512, iters = 1 << 15;
        float* d = nullptr;
        B200_CUDA(cudaMalloc(&d, sizeof(float) * blocks * threads));
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        double best = 0;
        for (int rep = 0; rep < 4; rep++)
            {
                cudaEventRecord(e0);
                fp32_peak_kernel<<<blocks, threads>>>(d, iters);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms = 0;
                cudaEventElapsedTime(&ms, e0, e1);
                const double tf = 2.0 * 8.0 * (double)iters * blocks * threads / (ms * 1e-3) / 1e12;
                if (rep > 0 && tf > best) best = tf;
            }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaFree(d);
        B200_CUDA(cudaGetLastError());
        *tflops = best;
        return GNSS_B200_OK;
    }

    const char* gnss_b200_last_error(void) { return b200::last_error_buf(); }
    const char* gnss_b200_version(void) { return "gnss_b200 0.1.0 (sm_100a)"; }

    int gnss_b200_device_count(void)
    {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
        return n;
    }

    int gnss_b200_host_register(void* ptr, size_t bytes)
    {
        B200_REQUIRE(ptr && bytes, GNSS_B200_ERR_INVALID, "host_register: bad arguments");
        B200_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
        return GNSS_B200_OK;
    }

    int gnss_b200_host_unregister(void* ptr)
    {
        B200_REQUIRE(ptr, GNSS_B200_ERR_INVALID, "host_unregister: null");
        B200_CUDA(cudaHostUnregister(ptr));
        return GNSS_B200_OK;
    }

    // pcps_acquisition::calculate_threshold (pcps_acquisition.cc:894-909)
    float gnss_b200_acq_threshold(float pfa, uint32_t effective_size, uint32_t n_doppler, uint32_t max_dwells, int bit_transition)
    {
        if (pfa <= 0.0f) return 0.0f;
        const int num_bins = (int)effective_size * (int)n_doppler;
        const double p = std::pow(1.0 - (double)pfa, 1.0 / (double)(float)num_bins);
        return (float)(2.0 * b200::gamma_p_inv(2.0 * (bit_transition ? 1.0 : (double)max_dwells), p));
    }
}
