// Gibbs read-count sampler on the device, with counter-based random numbers.
//
// Reference semantics restated (never its code): gibbsReadCountSampler (rpvg:src/path_abundance_estimator.cpp:116-212).
// Start from the EM abundances; per iteration: E-step posteriors; every read class splits its uint32 read count over its
// columns (ascending path index, noise last) by conditional binomials Bin(n_left, min(1, p / p_left)); every column draws
// Gamma(count + 1, 1); the draws are normalised (a Dirichlet sample); every gibbs_thin_its iterations the abundances times
// total_count are recorded, values below 1e-8 folded into the noise sample.
//
// The reference consumes one sequential mt19937 stream per cluster. Here every draw is addressed by
// Philox4x32-10(key = cluster seed, counter = (draw, iteration, read class or column, stream)), so the result does not
// depend on how read classes are spread over threads and a run is reproducible bit for bit; it cannot (and need not)
// reproduce the reference's stream: agreement is in distribution (tests/test_gibbs_gpu.py).
//   binomial: exact inversion — from 0 when n p < 12, from the mode outwards otherwise (expected steps ~ sqrt(n p q));
//   gamma (shape >= 1): Marsaglia-Tsang with Box-Muller normals. All fp64.
// One CTA per cluster; column counts are integer atomics (order-free, hence deterministic).

#include <vector>

#include "common.cuh"
#include "em_batch.cuh"

namespace {

constexpr int kGibbsThreads = 256;

struct Philox {

    uint2 key;
    uint4 counter;

    __device__ __forceinline__ static void round(uint4 & c, uint2 & k) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }

    __device__ __forceinline__ uint4 block() {
        uint4 c = counter;
        uint2 k = key;
        #pragma unroll
        for (int i = 0; i < 10; ++i) { round(c, k); }
        counter.x++;
        return c;
    }
};

struct Rng {

    Philox philox;
    uint4 buffer;
    int left;

    __device__ Rng(uint64_t seed, uint32_t iteration, uint32_t item, uint32_t stream) {
        philox.key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
        philox.counter = make_uint4(0, iteration, item, stream);
        left = 0;
    }

    // uniform in (0, 1), 53 random bits
    __device__ __forceinline__ double uniform() {
        if (left == 0) {
            buffer = philox.block();
            left = 2;
        }
        const uint64_t bits = (left == 2) ? ((static_cast<uint64_t>(buffer.x) << 32) | buffer.y) : ((static_cast<uint64_t>(buffer.z) << 32) | buffer.w);
        left--;
        return (static_cast<double>(bits >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    }

    __device__ __forceinline__ double normal() {
        const double u1 = uniform(), u2 = uniform();
        return sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
    }
};

__device__ uint32_t sampleBinomial(Rng & rng, const uint32_t n, double p) {

    if (n == 0 || !(p > 0)) { return 0; }
    if (p >= 1) { return n; }

    const bool flipped = p > 0.5;
    if (flipped) { p = 1 - p; }

    const double q = 1 - p;
    uint32_t x;

    if (n * p < 12) {

        // inversion from 0
        const double s = p / q;
        const double a = (n + 1.0) * s;
        double r = exp(n * log1p(-p));
        double u = rng.uniform();
        x = 0;
        while (u > r && x < n) {
            u -= r;
            ++x;
            r *= (a / x - s);
        }

    } else {

        // inversion from the mode outwards
        const uint32_t mode = min(n, static_cast<uint32_t>((n + 1.0) * p));
        const double f_mode = exp(lgamma(n + 1.0) - lgamma(mode + 1.0) - lgamma(n - mode + 1.0) + mode * log(p) + (n - mode) * log1p(-p));

        double u = rng.uniform();
        x = mode;

        if (u >= f_mode) {

            u -= f_mode;
            uint32_t lo = mode, hi = mode;
            double f_lo = f_mode, f_hi = f_mode;

            while (lo > 0 || hi < n) {

                if (lo > 0) {
                    f_lo *= (lo * q) / ((n - lo + 1.0) * p);
                    --lo;
                    if (u < f_lo) { x = lo; break; }
                    u -= f_lo;
                }
                if (hi < n) {
                    f_hi *= ((n - hi) * p) / ((hi + 1.0) * q);
                    ++hi;
                    if (u < f_hi) { x = hi; break; }
                    u -= f_hi;
                }
            }
        }
    }

    return flipped ? n - x : x;
}

// Gamma(shape, 1) for shape >= 1 (Marsaglia & Tsang 2000)
__device__ double sampleGamma(Rng & rng, const double shape) {

    const double d = shape - 1.0 / 3.0;
    const double c = 1.0 / sqrt(9.0 * d);

    for (int tries = 0; tries < 1000; ++tries) {

        const double x = rng.normal();
        double v = 1.0 + c * x;
        if (v <= 0) { continue; }
        v = v * v * v;

        const double u = rng.uniform();
        if (log(u) < 0.5 * x * x + d - d * v + d * log(v)) { return d * v; }
    }

    return d;
}

struct GibbsArgs {

    const uint32_t * cluster_list;     // clusters with samples to draw
    const uint64_t * read_offset;
    const uint64_t * path_offset;
    const uint64_t * row_ptr;          // the batch's sorted CSR
    const uint32_t * col;
    const double * val;
    const double * noise;
    const double * cnt;
    const double * em_abundances;      // EM results (read counts)
    const double * em_noise_count;
    const double * em_total_count;
    const uint32_t * num_samples;      // [n_clusters]
    const uint64_t * seeds;            // [n_clusters]
    const uint64_t * sample_offset;    // [n_clusters] first sample of the cluster in noise_samples
    const uint64_t * abundance_offset; // [n_clusters] first value of the cluster in abundance_samples
    uint32_t thin_its;
    double * state;                    // [n_paths + n_clusters] relative abundances per cluster (noise last)
    uint32_t * column_counts;          // [n_paths + n_clusters]
    double * abundance_samples;
    double * noise_samples;
};

__global__ void __launch_bounds__(kGibbsThreads) gibbsKernel(const GibbsArgs g) {

    __shared__ double red[32];

    const int tid = threadIdx.x;
    const uint32_t c = g.cluster_list[blockIdx.x];

    const uint64_t r0 = g.read_offset[c];
    const uint32_t M = static_cast<uint32_t>(g.read_offset[c + 1] - r0);
    const uint64_t p0 = g.path_offset[c];
    const uint32_t N = static_cast<uint32_t>(g.path_offset[c + 1] - p0);

    const uint32_t samples = g.num_samples[c];
    const uint64_t seed = g.seeds[c];
    const double total = g.em_total_count[c];

    double * a = g.state + p0 + c;
    uint32_t * counts = g.column_counts + p0 + c;
    double * out_abundance = g.abundance_samples + g.abundance_offset[c];
    double * out_noise = g.noise_samples + g.sample_offset[c];

    if (M == 0 || !(total > 0)) {
        for (uint64_t k = tid; k < static_cast<uint64_t>(samples) * N; k += kGibbsThreads) { out_abundance[k] = 0; }
        for (uint32_t s = tid; s < samples; s += kGibbsThreads) { out_noise[s] = 0; }
        return;
    }

    // rpvg:src/path_abundance_estimator.cpp:129-136
    for (uint32_t j = tid; j < N; j += kGibbsThreads) { a[j] = g.em_abundances[p0 + j] / total; }
    if (tid == 0) { a[N] = g.em_noise_count[c] / total; }
    __syncthreads();

    const uint32_t iterations = samples * g.thin_its;

    for (uint32_t it = 1; it <= iterations; ++it) {

        for (uint32_t j = tid; j <= N; j += kGibbsThreads) { counts[j] = 0; }
        __syncthreads();

        const double a_noise = a[N];

        for (uint32_t i = tid; i < M; i += kGibbsThreads) {

            const uint64_t e0 = g.row_ptr[r0 + i], e1 = g.row_ptr[r0 + i + 1];
            const double nz = g.noise[r0 + i];

            // posteriors p_ij a_j / sum (left-to-right sum, noise last)
            double row_sum = 0;
            for (uint64_t e = e0; e < e1; ++e) { row_sum += g.val[e] * a[g.col[e]]; }
            row_sum += nz * a_noise;

            uint32_t reads_left = static_cast<uint32_t>(g.cnt[r0 + i]);
            double prob_left = 1;

            Rng rng(seed, it, i, 0);

            for (uint64_t e = e0; e < e1 && reads_left > 0; ++e) {

                const double cur_prob = (g.val[e] * a[g.col[e]]) / row_sum;

                if (cur_prob > 0) {
                    const uint32_t drawn = sampleBinomial(rng, reads_left, fmin(1.0, cur_prob / prob_left));
                    if (drawn) { atomicAdd(&counts[g.col[e]], drawn); }
                    reads_left -= drawn;
                }
                prob_left -= cur_prob;
            }

            // the noise column is last and takes what is left (the reference draws Bin(n_left, min(1, ~1)) and asserts
            // that nothing remains)
            if (reads_left > 0) { atomicAdd(&counts[N], reads_left); }
        }

        __syncthreads();

        // Dirichlet draw: Gamma(count + gamma(=1), 1) per column, normalised
        double partial = 0;
        for (uint32_t j = tid; j <= N; j += kGibbsThreads) {
            Rng rng(seed, it, j, 1);
            const double x = sampleGamma(rng, counts[j] + 1.0);
            a[j] = x;
            partial += x;
        }

        const double gamma_sum = blockSum(partial, red);

        for (uint32_t j = tid; j <= N; j += kGibbsThreads) { a[j] = a[j] / gamma_sum; }
        __syncthreads();

        if (it % g.thin_its == 0) {

            const uint32_t s = it / g.thin_its - 1;
            double noise_part = 0;

            for (uint32_t j = tid; j < N; j += kGibbsThreads) {
                const double aj = a[j];
                if (aj < 1e-8) {   // min_gibbs_abundance, rpvg:src/path_abundance_estimator.cpp:14
                    noise_part += aj * total;
                    out_abundance[static_cast<uint64_t>(s) * N + j] = 0;
                } else {
                    out_abundance[static_cast<uint64_t>(s) * N + j] = aj * total;
                }
            }

            const double low_sum = blockSum(noise_part, red);
            if (tid == 0) { out_noise[s] = low_sum + a[N] * total; }
        }
    }
}

}

// Draws num_samples[c] Gibbs samples for every cluster of a batch whose EM has run. Host pointers out.
int rpvgBatchGibbs(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, const uint32_t * num_samples, uint32_t thin_its, const uint64_t * seeds, double * abundance_samples,
                   double * noise_samples) {

    if (!b->have_results) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "the EM has not run on this batch"); }
    if (thin_its == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "gibbs thin_its must be > 0"); }

    const uint32_t C = b->n_clusters;

    std::vector<uint32_t> cluster_list;
    std::vector<uint64_t> sample_offset(C + 1, 0), abundance_offset(C + 1, 0);

    for (uint32_t c = 0; c < C; ++c) {
        if (num_samples[c] > 0) { cluster_list.push_back(c); }
        sample_offset[c + 1] = sample_offset[c] + num_samples[c];
        abundance_offset[c + 1] = abundance_offset[c] + static_cast<uint64_t>(num_samples[c]) * (b->h_path_offset[c + 1] - b->h_path_offset[c]);
    }

    if (cluster_list.empty()) { return RPVG_B200_OK; }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    auto pad = [](size_t n) { return (n + 255) & ~size_t(255); };
    const size_t bytes = pad(4 * cluster_list.size()) + pad(4 * C) + 3 * pad(8 * (C + 1)) + pad(4 * (b->n_paths + C)) + pad(8 * abundance_offset[C]) + pad(8 * sample_offset[C]) + 1024;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t n) { char * q = base + used; used += pad(n); return q; };

    uint32_t * d_list = reinterpret_cast<uint32_t *>(take(4 * cluster_list.size()));
    uint32_t * d_num_samples = reinterpret_cast<uint32_t *>(take(4 * C));
    uint64_t * d_seeds = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_sample_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_abundance_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint32_t * d_counts = reinterpret_cast<uint32_t *>(take(4 * (b->n_paths + C)));
    double * d_abundance_samples = reinterpret_cast<double *>(take(8 * abundance_offset[C]));
    double * d_noise_samples = reinterpret_cast<double *>(take(8 * sample_offset[C]));

    cudaError_t err = cudaSuccess;
    #define RPVG_STEP(call) if (err == cudaSuccess) { err = (call); }

    RPVG_STEP(cudaMemcpyAsync(d_list, cluster_list.data(), 4 * cluster_list.size(), cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_num_samples, num_samples, 4 * C, cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_seeds, seeds, 8 * C, cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_sample_offset, sample_offset.data(), 8 * (C + 1), cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_abundance_offset, abundance_offset.data(), 8 * (C + 1), cudaMemcpyHostToDevice, s))

    if (err == cudaSuccess) {

        GibbsArgs g;
        g.cluster_list = d_list;
        g.read_offset = b->d_read_offset;
        g.path_offset = b->d_path_offset;
        g.row_ptr = b->d_sorted_row_ptr;
        g.col = b->d_sorted_col;
        g.val = b->d_val;
        g.noise = b->d_sorted_noise;
        g.cnt = b->d_cnt;
        g.em_abundances = b->d_abundances;
        g.em_noise_count = b->d_noise_count;
        g.em_total_count = b->d_total_count;
        g.num_samples = d_num_samples;
        g.seeds = d_seeds;
        g.sample_offset = d_sample_offset;
        g.abundance_offset = d_abundance_offset;
        g.thin_its = thin_its;
        g.state = b->d_abund;
        g.column_counts = d_counts;
        g.abundance_samples = d_abundance_samples;
        g.noise_samples = d_noise_samples;

        gibbsKernel<<<static_cast<unsigned>(cluster_list.size()), kGibbsThreads, 0, s>>>(g);
        ctx->launches++;
        err = cudaGetLastError();
    }

    RPVG_STEP(cudaMemcpyAsync(abundance_samples, d_abundance_samples, 8 * abundance_offset[C], cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaMemcpyAsync(noise_samples, d_noise_samples, 8 * sample_offset[C], cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaStreamSynchronize(s))
    #undef RPVG_STEP

    cudaFreeAsync(base, s);

    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("gibbs read-count sampler: ") + cudaGetErrorString(err)); }
    return RPVG_B200_OK;
}

extern "C" int rpvg_b200_batch_gibbs(rpvg_b200_ctx * ctx, rpvg_b200_batch * batch, const uint32_t * num_samples, uint32_t thin_its, const uint64_t * seeds,
                                     double * abundance_samples, double * noise_samples) {

    if (!ctx || !batch || !num_samples || !seeds || !abundance_samples || !noise_samples) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    return rpvgBatchGibbs(ctx, batch, num_samples, thin_its, seeds, abundance_samples, noise_samples);
}
