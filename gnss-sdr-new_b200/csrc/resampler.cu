// The "acquisition resampler" (SURVEY 8.6 N2): the decimating low-pass FIR that GNSSFlowgraph puts in front of the
// acquisition blocks when GNSS-SDR.use_acquisition_resampler is set (gnss_flowgraph.cc:1063-1118: decimation from
// Acq_Conf::ConfigureAutomaticResampler, acq_conf.cc:93-111; taps = gr::filter::firdes::low_pass(1.0, fs,
// acq_fs/2.1, acq_fs/2); block = gr::filter::fir_filter_ccf::make(decimation, taps)).
//
// Both pieces are GNU Radio code, a dependency that is not part of the reference tree; what is restated here is their
// published algorithm (gr-filter/lib/firdes.cc: windowed-sinc design with a Hamming window, tap count from the
// window's 53 dB attenuation; fir_filter_ccf: y[m] = sum_j taps[j] * x[m*D - j] with ntaps-1 items of history, zeros at
// the start of the stream). The decimated block stays in device memory, where gnss_b200_acq_dwell_device_async reads
// it, so the raw-rate samples cross PCIe once and the decimated ones never.
#include "common.cuh"
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

using namespace b200;

namespace
{
constexpr uint32_t kMaxTaps = 4096;
constexpr int kThreads = 256;

// one thread per output item; taps staged in shared memory, reversed so that the input is walked forwards
__global__ void __launch_bounds__(kThreads) fir_decimate_kernel(const float2* __restrict__ buf, const float* __restrict__ taps, uint32_t n_taps,
    uint32_t decimation, uint32_t n_out, float2* __restrict__ out)
{
    extern __shared__ float s_taps[];
    for (uint32_t i = threadIdx.x; i < n_taps; i += kThreads) s_taps[i] = taps[n_taps - 1 - i];
    __syncthreads();
    const uint32_t m = blockIdx.x * kThreads + threadIdx.x;
    if (m >= n_out) return;
    // buf = [ntaps-1 items of history | this call's items]: item x[m*D - j] sits at (ntaps-1) + m*D - j
    const float2* x = buf + (size_t)m * decimation;
    float re = 0.f, im = 0.f;
    for (uint32_t k = 0; k < n_taps; k++)
        {
            const float2 v = __ldg(x + k);
            re = fmaf(v.x, s_taps[k], re);
            im = fmaf(v.y, s_taps[k], im);
        }
    out[m] = make_float2(re, im);
}
}  // namespace

struct gnss_b200_resampler
{
    std::mutex mu;
    int device = 0;
    uint32_t decimation = 1, n_taps = 0, max_in = 0;
    cudaStream_t stream = nullptr;
    DevBuf d_taps, d_buf, d_out;
    PinBuf h_in, h_out;
};

extern "C"
{
    int gnss_b200_firdes_low_pass(double gain, double sampling_freq, double cutoff_freq, double transition_width, uint32_t max_taps, float* taps,
        uint32_t* n_taps)
    {
        B200_REQUIRE(taps && n_taps, GNSS_B200_ERR_INVALID, "firdes_low_pass: null output");
        B200_REQUIRE(sampling_freq > 0.0 && cutoff_freq > 0.0 && cutoff_freq <= sampling_freq / 2 && transition_width > 0.0, GNSS_B200_ERR_INVALID,
            "firdes_low_pass: fs=%g cutoff=%g transition=%g", sampling_freq, cutoff_freq, transition_width);
        // compute_ntaps: Hamming window, max attenuation 53 dB; odd length
        int ntaps = (int)(53.0 * sampling_freq / (22.0 * transition_width));
        if ((ntaps & 1) == 0) ntaps++;
        B200_REQUIRE(ntaps >= 1 && (uint32_t)ntaps <= max_taps && (uint32_t)ntaps <= kMaxTaps, GNSS_B200_ERR_INVALID, "firdes_low_pass: %d taps do not fit (%u)", ntaps,
            max_taps);
        const double pi = 3.14159265358979323846;
        const int M = (ntaps - 1) / 2;
        const double fwT0 = 2 * pi * cutoff_freq / sampling_freq;
        for (int n = -M; n <= M; n++)
            {
                const float w = ntaps == 1 ? 1.0f : (float)(0.54 - 0.46 * std::cos((2 * pi * (n + M)) / (ntaps - 1)));  // window::hamming
                if (n == 0)
                    taps[n + M] = (float)(fwT0 / pi * w);
                else
                    taps[n + M] = (float)(std::sin(n * fwT0) / (n * pi) * w);
            }
        // unity gain at DC
        double fmax = taps[0 + M];
        for (int n = 1; n <= M; n++) fmax += 2 * taps[n + M];
        gain /= fmax;
        for (int i = 0; i < ntaps; i++) taps[i] = (float)(taps[i] * gain);
        *n_taps = (uint32_t)ntaps;
        return GNSS_B200_OK;
    }

    int gnss_b200_resampler_create(uint32_t decimation, uint32_t n_taps, const float* taps, uint32_t max_input_items, int32_t device,
        gnss_b200_resampler** out)
    {
        B200_REQUIRE(out && taps, GNSS_B200_ERR_INVALID, "resampler_create: null argument");
        *out = nullptr;
        B200_REQUIRE(decimation >= 1 && n_taps >= 1 && n_taps <= kMaxTaps && max_input_items >= decimation, GNSS_B200_ERR_INVALID,
            "resampler_create: decimation=%u n_taps=%u max_input_items=%u", decimation, n_taps, max_input_items);
        int rc = select_device(device);
        if (rc) return rc;
        auto* h = new gnss_b200_resampler();
        h->device = device;
        h->decimation = decimation;
        h->n_taps = n_taps;
        h->max_in = max_input_items;
        auto cleanup = [&](int code) {
            gnss_b200_resampler_destroy(h);
            return code;
        };
        if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) return cleanup(fail(GNSS_B200_ERR_CUDA, "resampler_create: stream"));
        const size_t buf_items = (size_t)n_taps - 1 + max_input_items;
        if (h->d_taps.reserve(sizeof(float) * n_taps) != cudaSuccess || h->d_buf.reserve(sizeof(float2) * buf_items) != cudaSuccess ||
            h->d_out.reserve(sizeof(float2) * (max_input_items / decimation)) != cudaSuccess || h->h_in.reserve(sizeof(float2) * max_input_items) != cudaSuccess ||
            h->h_out.reserve(sizeof(float2) * (max_input_items / decimation)) != cudaSuccess)
            return cleanup(fail(GNSS_B200_ERR_CUDA, "resampler_create: out of memory"));
        if (cudaMemcpy(h->d_taps.p, taps, sizeof(float) * n_taps, cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaMemset(h->d_buf.p, 0, sizeof(float2) * buf_items) != cudaSuccess)
            return cleanup(fail(GNSS_B200_ERR_CUDA, "resampler_create: upload"));
        *out = h;
        return GNSS_B200_OK;
    }

    int gnss_b200_resampler_destroy(gnss_b200_resampler* h)
    {
        if (!h) return GNSS_B200_OK;
        cudaSetDevice(h->device);
        if (h->stream) cudaStreamSynchronize(h->stream);
        h->d_taps.release();
        h->d_buf.release();
        h->d_out.release();
        h->h_in.release();
        h->h_out.release();
        if (h->stream) cudaStreamDestroy(h->stream);
        delete h;
        return GNSS_B200_OK;
    }

    int gnss_b200_resampler_reset(gnss_b200_resampler* h)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "resampler_reset: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        B200_CUDA(cudaMemsetAsync(h->d_buf.p, 0, sizeof(float2) * (h->n_taps - 1), h->stream));
        B200_CUDA(cudaStreamSynchronize(h->stream));
        return GNSS_B200_OK;
    }

    int gnss_b200_resampler_work(gnss_b200_resampler* h, const float* in_c64, uint32_t n_input_items, float* out_c64, const void** out_device)
    {
        B200_REQUIRE(h && in_c64, GNSS_B200_ERR_INVALID, "resampler_work: null argument");
        B200_REQUIRE(n_input_items >= h->decimation && n_input_items <= h->max_in && n_input_items % h->decimation == 0, GNSS_B200_ERR_INVALID,
            "resampler_work: %u input items (a multiple of the decimation %u, at most %u)", n_input_items, h->decimation, h->max_in);
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        const uint32_t hist = h->n_taps - 1, n_out = n_input_items / h->decimation;
        float2* buf = h->d_buf.as<float2>();
        std::memcpy(h->h_in.p, in_c64, sizeof(float2) * n_input_items);
        B200_CUDA(cudaMemcpyAsync(buf + hist, h->h_in.p, sizeof(float2) * n_input_items, cudaMemcpyHostToDevice, h->stream));
        fir_decimate_kernel<<<(n_out + kThreads - 1) / kThreads, kThreads, sizeof(float) * h->n_taps, h->stream>>>(buf, h->d_taps.as<float>(), h->n_taps,
            h->decimation, n_out, h->d_out.as<float2>());
        B200_CUDA(cudaGetLastError());
        if (out_c64) B200_CUDA(cudaMemcpyAsync(h->h_out.p, h->d_out.p, sizeof(float2) * n_out, cudaMemcpyDeviceToHost, h->stream));
        // the last ntaps-1 items become the history of the next call
        if (hist > 0)
            {
                if (n_input_items >= hist)
                    B200_CUDA(cudaMemcpyAsync(buf, buf + n_input_items, sizeof(float2) * hist, cudaMemcpyDeviceToDevice, h->stream));
                else
                    {
                        // shift the old history left by n_input_items, one sync copy each way (short blocks only)
                        std::vector<float2> tmp(hist + n_input_items);
                        B200_CUDA(cudaStreamSynchronize(h->stream));
                        B200_CUDA(cudaMemcpy(tmp.data(), buf, sizeof(float2) * tmp.size(), cudaMemcpyDeviceToHost));
                        B200_CUDA(cudaMemcpy(buf, tmp.data() + n_input_items, sizeof(float2) * hist, cudaMemcpyHostToDevice));
                    }
            }
        B200_CUDA(cudaStreamSynchronize(h->stream));
        if (out_c64) std::memcpy(out_c64, h->h_out.p, sizeof(float2) * n_out);
        if (out_device) *out_device = h->d_out.p;
        return GNSS_B200_OK;
    }
}
