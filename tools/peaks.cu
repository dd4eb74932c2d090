// tools/peaks.cu — measured issue / pipe / cache peaks of the box's B200, the denominators of the
// `roofline` object bench.py prints for the issue-bound ray kernels ("Peak
// issue/FP64/FP32 rates must be measured on the box with an FMA microbenchmark").
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/peaks tools/peaks.cu && tools/peaks > profiles/r2_peaks.json
//
// Every kernel is a register-resident loop of independent dependency chains (8 per thread) at full
// occupancy, timed with CUDA events after a warm-up launch; the figure is the best of 5 launches.
//   * fp64_fma / fp64_mul_add  DFMA and the unfused DMUL+DADD pair (librptx is built -fmad=false)
//   * fp32_fma                 FFMA
//   * int_imad / int_lop       integer multiply-add, logic ops
//   * issue_mixed              FFMA + IADD3 + LOP3 interleaved (two pipes): warp instructions per second
//                              the four schedulers of an SM can issue, the roof of `achieved` for the
//                              ray kernels (thread-instructions/s = 32 x that, when every lane is live)
//   * l2_read                  grid-stride float4 reads of a 48 MB buffer (L2-resident after the first pass)
//   * l1_read                  every block re-reads its own 32 KB window (L1-resident)
//   * hbm_copy                 float4 copy of 2 GiB (read + write bytes)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int kChains = 8;
constexpr int kInner = 64;  // unrolled instructions per chain and loop trip

template <int OP>
__global__ void __launch_bounds__(256) k_alu(int trips, double seed, unsigned m1, unsigned m2, double* out) {
    double d[kChains];
    float f[kChains];
    unsigned u[kChains];
#pragma unroll
    for (int c = 0; c < kChains; c++) { d[c] = seed + c; f[c] = (float)seed + c; u[c] = (unsigned)(seed * 17.0) + c + threadIdx.x; }
    const double da = 1.0000001, db = 1e-9;
    const float fa = 1.0000001f, fb = 1e-9f;
    for (int t = 0; t < trips; t++) {
#pragma unroll
        for (int i = 0; i < kInner; i++) {
#pragma unroll
            for (int c = 0; c < kChains; c++) {
                if (OP == 0) d[c] = fma(d[c], da, db);
                if (OP == 1) d[c] = __dadd_rn(__dmul_rn(d[c], da), db);
                if (OP == 2) f[c] = fmaf(f[c], fa, fb);
                // m1 / m2 are run-time values and the integer chains read their neighbour, so that the
                // compiler cannot collapse 64 steps into one
                if (OP == 3) u[c] = u[c] * m1 + u[(c + 1) & 7];
                if (OP == 4) u[c] = u[c] ^ (u[(c + 1) & 7] & m2);
                if (OP == 5) {  // two pipes interleaved: fp32 FMA, integer add / logic
                    if ((c & 1) == 0) f[c] = fmaf(f[c], fa, fb);
                    else if ((c & 3) == 1) u[c] = u[c] + u[(c + 2) & 7];
                    else u[c] = u[c] ^ (u[(c + 2) & 7] & m2);
                }
            }
        }
    }
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < kChains; c++) s += d[c] + (double)f[c] + (double)u[c];
    if (s == 12345.678) out[0] = s;  // never true: keeps the chains alive
}

__global__ void __launch_bounds__(256) k_read(const float4* __restrict__ p, size_t n, int passes, float* out) {
    float acc = 0.f;
    for (int r = 0; r < passes; r++)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
            const float4 v = __ldcg(p + i);  // L2 only
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 12345.678f) out[0] = acc;
}

__global__ void __launch_bounds__(256) k_read_l1(const float4* __restrict__ p, int window, int passes, float* out) {
    const float4* base = p + (size_t)blockIdx.x * window;
    float acc = 0.f;
    for (int r = 0; r < passes; r++)
        for (int i = threadIdx.x; i < window; i += blockDim.x) {
            const float4 v = __ldca(base + i);
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 12345.678f) out[0] = acc;
}

__global__ void __launch_bounds__(256) k_copy(const float4* __restrict__ src, float4* __restrict__ dst, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

template <class F>
float best_ms(F launch, int reps = 5) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    launch();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        CK(cudaEventRecord(a));
        launch();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int clock_khz = 0;
    CK(cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0));
    double* d_out;
    CK(cudaMalloc(&d_out, 64));
    const int blocks = sms * 8, threads = 256, trips = 256;
    const unsigned m1 = (unsigned)(clock_khz | 1), m2 = 0x7f4a7c15u ^ (unsigned)sms;  // not compile-time constants
    const double insts = (double)blocks * threads * trips * kInner * kChains;  // thread instructions (OP 1: two per step)
    printf("{\n \"gpu\": \"%s\", \"sms\": %d, \"sm_clock_max_mhz\": %.0f,\n", prop.name, sms, clock_khz / 1e3);
    const char* names[6] = {"fp64_fma", "fp64_mul_add", "fp32_fma", "int_imad", "int_lop", "issue_mixed"};
    for (int op = 0; op < 6; op++) {
        float ms = 0;
        switch (op) {
            case 0: ms = best_ms([&] { k_alu<0><<<blocks, threads>>>(trips, 1.0, m1, m2, d_out); }); break;
            case 1: ms = best_ms([&] { k_alu<1><<<blocks, threads>>>(trips, 1.0, m1, m2, d_out); }); break;
            case 2: ms = best_ms([&] { k_alu<2><<<blocks, threads>>>(trips, 1.0, m1, m2, d_out); }); break;
            case 3: ms = best_ms([&] { k_alu<3><<<blocks, threads>>>(trips, 1.0, m1, m2, d_out); }); break;
            case 4: ms = best_ms([&] { k_alu<4><<<blocks, threads>>>(trips, 1.0, m1, m2, d_out); }); break;
            default: ms = best_ms([&] { k_alu<5><<<blocks, threads>>>(trips, 1.0, m1, m2, d_out); }); break;
        }
        const double per_step = op == 1 ? 2.0 : 1.0;  // DMUL + DADD
        const double tinst = insts * per_step / (ms * 1e-3);
        printf(" \"%s\": {\"ms\": %.4f, \"thread_inst_per_s\": %.4e, \"warp_inst_per_s\": %.4e, \"warp_inst_per_clk_per_sm\": %.3f},\n", names[op], ms,
               tinst, tinst / 32.0, tinst / 32.0 / sms / (clock_khz * 1e3));
    }
    {
        const size_t bytes = 48ull << 20, n = bytes / 16;
        float4* buf;
        float* fo;
        CK(cudaMalloc(&buf, bytes)); CK(cudaMalloc(&fo, 64));
        CK(cudaMemset(buf, 0, bytes));
        const int passes = 20;
        const float ms = best_ms([&] { k_read<<<sms * 16, 256>>>(buf, n, passes, fo); });
        printf(" \"l2_read\": {\"ms\": %.4f, \"GBps\": %.1f, \"bytes\": %zu},\n", ms, (double)bytes * passes / (ms * 1e-3) / 1e9, bytes);
        const int window = 2048;  // float4 per block = 32 KB
        const float ms1 = best_ms([&] { k_read_l1<<<sms * 4, 256>>>(buf, window, 200, fo); });
        printf(" \"l1_read\": {\"ms\": %.4f, \"GBps\": %.1f},\n", ms1, (double)sms * 4 * window * 16 * 200 / (ms1 * 1e-3) / 1e9);
        CK(cudaFree(buf)); CK(cudaFree(fo));
    }
    {
        const size_t bytes = 2ull << 30, n = bytes / 16;
        float4 *a, *b;
        CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes));
        CK(cudaMemset(a, 1, bytes));
        const float ms = best_ms([&] { k_copy<<<sms * 16, 256>>>(a, b, n); });
        printf(" \"hbm_copy\": {\"ms\": %.4f, \"GBps\": %.1f}\n}\n", ms, 2.0 * bytes / (ms * 1e-3) / 1e9);
        CK(cudaFree(a)); CK(cudaFree(b));
    }
    return 0;
}
