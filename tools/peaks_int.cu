// Measured integer / byte-SIMD / shared-memory issue peaks of the B200 (sm_100a).
//
// The roofs that bound FAST, rBRIEF, Hamming and the fixed-point pyramid / blur arithmetic are instruction
// issue rates, not HBM bandwidth.  This tool measures them: every test is a dependent-free unrolled loop of
// one instruction (or of a two-instruction mix, to see which instructions share a pipe), run by one CTA of
// 1024 threads per SM for long enough to reach steady state.  Reported per test:
//   lanes_per_clk_per_sm  thread-level operations per SM clock per SM, from clock64() inside the kernel
//                         (independent of the clock the GPU happens to run at)
//   gops                  thread-level operations per second, from CUDA events (at the clock of the run)
// Output: one JSON object on stdout (profiles/MEASURED_INT_PEAKS.json is this output, verbatim).
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o peaks_int peaks_int.cu && ./peaks_int
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <string>
#include <algorithm>

#define CHAINS 8      // independent dependency chains per thread
#define UNROLL 16     // operations per chain per loop iteration
#define THREADS 1024

struct Times { long long t0, t1; };

// Each OP(x, a, b) must be ONE SASS instruction of the class under test (checked with cuobjdump -sass: 1920 of the
// instruction per kernel after unrolling).  Operands a, b come from global memory so that
// nothing folds at compile time.
#define DEFINE_TEST(NAME, BODY)                                                                                  \
    __global__ void __launch_bounds__(THREADS) NAME(const unsigned* __restrict__ in, unsigned* __restrict__ out,  \
                                                   Times* __restrict__ tm, const int iters) {                    \
        unsigned x[CHAINS];                                                                                      \
        const unsigned a = in[threadIdx.x & 31], b = in[32 + (threadIdx.x & 31)]; (void)b;                      \
        _Pragma("unroll") for (int c = 0; c < CHAINS; ++c) x[c] = in[64 + c] ^ threadIdx.x;                      \
        __syncthreads();                                                                                         \
        const long long t0 = clock64();                                                                          \
        for (int i = 0; i < iters; ++i) {                                                                        \
            _Pragma("unroll") for (int u = 0; u < UNROLL; ++u) {                                                 \
                _Pragma("unroll") for (int c = 0; c < CHAINS; ++c) { BODY }                                      \
            }                                                                                                    \
        }                                                                                                        \
        const long long t1 = clock64();                                                                          \
        unsigned r = 0;                                                                                          \
        _Pragma("unroll") for (int c = 0; c < CHAINS; ++c) r ^= x[c];                                            \
        if (r == 0x12345678u) out[threadIdx.x] = r;                                                              \
        if (threadIdx.x == 0) { tm[blockIdx.x].t0 = t0; tm[blockIdx.x].t1 = t1; }                                \
    }

__device__ __forceinline__ unsigned op_popc(unsigned x) { unsigned r; asm("popc.b32 %0, %1;" : "=r"(r) : "r"(x)); return r; }
__device__ __forceinline__ unsigned op_lop3(unsigned x, unsigned a, unsigned b) { unsigned r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(x), "r"(a), "r"(b)); return r; }
__device__ __forceinline__ unsigned op_iadd3(unsigned x, unsigned a, unsigned b) { unsigned r; asm("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, t, %3; }" : "=r"(r) : "r"(x), "r"(a), "r"(b)); return r; }
__device__ __forceinline__ unsigned op_imad(unsigned x, unsigned a, unsigned b) { unsigned r; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(a), "r"(b)); return r; }
__device__ __forceinline__ unsigned op_shf(unsigned x, unsigned a, unsigned b) { unsigned r; asm("shf.l.wrap.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(a), "r"(b)); return r; }
__device__ __forceinline__ unsigned op_prmt(unsigned x, unsigned a, unsigned b) { unsigned r; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(a), "r"(b)); return r; }

// one instruction per operation
DEFINE_TEST(t_popc,      x[c] = op_popc(x[c]);)
DEFINE_TEST(t_lop3,      x[c] = op_lop3(x[c], a, b);)
DEFINE_TEST(t_iadd3,     x[c] = op_iadd3(x[c], a, b);)
DEFINE_TEST(t_imad,      x[c] = op_imad(x[c], a, b);)
DEFINE_TEST(t_shf,       x[c] = op_shf(x[c], a, b);)
DEFINE_TEST(t_prmt,      x[c] = op_prmt(x[c], a, b);)
DEFINE_TEST(t_vimnmx3,   x[c] = __vimax3_s16x2(x[c], a, b);)
DEFINE_TEST(t_vabsdiff4, x[c] = __vabsdiffu4(x[c], a);)
DEFINE_TEST(t_dp4a,      x[c] = __dp4a(a, b, x[c]);)
DEFINE_TEST(t_dp2a,      x[c] = __dp2a_lo(a, b, x[c]);)
// two-instruction mixes (each counts as 2 operations): a sum close to the two single rates means different pipes
DEFINE_TEST(t_xor_popc,      x[c] = op_popc(x[c] ^ a);)                      // LOP3 + POPC (one Hamming word without the add)
DEFINE_TEST(t_lop3_imad,     x[c] = op_imad(op_lop3(x[c], a, b), a, b);)     // alu pipe + fma pipe
DEFINE_TEST(t_dp4a_lop3,     x[c] = op_lop3(__dp4a(a, b, x[c]), a, b);)
DEFINE_TEST(t_dp4a_imad,     x[c] = op_imad(__dp4a(a, b, x[c]), a, b);)
DEFINE_TEST(t_prmt_dp4a,     x[c] = __dp4a(op_prmt(x[c], a, b), b, x[c]);)
DEFINE_TEST(t_vimnmx3_lop3,  x[c] = op_lop3(__vimax3_s16x2(x[c], a, b), a, b);)
DEFINE_TEST(t_vimnmx3_imad,  x[c] = op_imad(__vimax3_s16x2(x[c], a, b), a, b);)
// half2 min / max (HMNMX2): could carry half of the FAST arc network (pixel values are exact in fp16) IF it issued on another pipe
__device__ __forceinline__ unsigned op_hmax2(unsigned x, unsigned a) { unsigned r; asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(a)); return r; }
DEFINE_TEST(t_hmnmx2,        x[c] = op_hmax2(x[c], a);)
DEFINE_TEST(t_vimnmx3_hmnmx2, x[c] = op_hmax2(__vimax3_s16x2(x[c], a, b), a);)
DEFINE_TEST(t_hmnmx2_imad,   x[c] = op_imad(op_hmax2(x[c], a), a, b);)
DEFINE_TEST(t_vhmnmx,        x[c] = op_hmax2(op_hmax2(x[c], a), b);)        // ptxas fuses the pair into one three-input VHMNMX (counted as ONE operation)
DEFINE_TEST(t_hmnmx2_lop3,   x[c] = op_lop3(op_hmax2(x[c], a), a, b);)

// Hamming distance of 256-bit descriptors exactly as the matcher kernels do it (8 x (LOP3 + POPC + IADD));
// operations counted = POPCs.
__global__ void __launch_bounds__(THREADS) t_hamming256(const unsigned* __restrict__ in, unsigned* __restrict__ out, Times* __restrict__ tm, const int iters) {
    unsigned q[8], acc[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 8; ++k) q[k] = in[k] ^ threadIdx.x;
    __shared__ unsigned s[8 * 64];
    for (int i = threadIdx.x; i < 8 * 64; i += THREADS) s[i] = in[i & 127] * (i + 1);
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const uint4 b0 = reinterpret_cast<const uint4*>(s)[(2 * (u + i)) & 127], b1 = reinterpret_cast<const uint4*>(s)[(2 * (u + i) + 1) & 127];
            acc[u & 3] += __popc(q[0] ^ b0.x) + __popc(q[1] ^ b0.y) + __popc(q[2] ^ b0.z) + __popc(q[3] ^ b0.w) +
                          __popc(q[4] ^ b1.x) + __popc(q[5] ^ b1.y) + __popc(q[6] ^ b1.z) + __popc(q[7] ^ b1.w);
        }
    }
    const long long t1 = clock64();
    const unsigned r = acc[0] ^ acc[1] ^ acc[2] ^ acc[3];
    if (r == 0x12345678u) out[threadIdx.x] = r;
    if (threadIdx.x == 0) { tm[blockIdx.x].t0 = t0; tm[blockIdx.x].t1 = t1; }
}

// shared-memory loads: conflict-free (consecutive words), broadcast (one word per warp), and byte loads
template <int MODE>
__global__ void __launch_bounds__(THREADS) t_lds(const unsigned* __restrict__ in, unsigned* __restrict__ out, Times* __restrict__ tm, const int iters) {
    __shared__ unsigned s[6144];
    for (int i = threadIdx.x; i < 6144; i += THREADS) s[i] = in[i & 127] + i;
    __syncthreads();
    unsigned acc[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) acc[c] = 0;
    const int warp = threadIdx.x >> 5;
    // every load below is base + immediate offset: no address arithmetic in the loop
    const volatile unsigned* vs = s + (MODE == 1 ? warp * 32 : threadIdx.x);
    const volatile unsigned char* vb = reinterpret_cast<const volatile unsigned char*>(s) + threadIdx.x * 4 + (threadIdx.x & 3);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
            for (int c = 0; c < CHAINS; ++c) {
                if (MODE == 2) acc[c] += vb[(u * CHAINS + c) * 128];   // LDS.U8, one byte per lane, distinct banks
                else acc[c] += vs[(u * CHAINS + c) * 32];
            }
        }
    }
    const long long t1 = clock64();
    unsigned r = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) r ^= acc[c];
    if (r == 0x12345678u) out[threadIdx.x] = r;
    if (threadIdx.x == 0) { tm[blockIdx.x].t0 = t0; tm[blockIdx.x].t1 = t1; }
}

typedef void (*kern_t)(const unsigned*, unsigned*, Times*, int);
struct Test { const char* name; kern_t k; double ops_per_thread_iter; const char* what; };

int main(int argc, char** argv) {
    int dev = 0;
    cudaSetDevice(dev);
    cudaDeviceProp pr;
    if (cudaGetDeviceProperties(&pr, dev) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 1; }
    const int nsm = pr.multiProcessorCount;
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
    unsigned h_in[4096];
    srand(7);
    for (int i = 0; i < 4096; ++i) h_in[i] = ((unsigned)rand() << 16) ^ (unsigned)rand();
    unsigned *d_in, *d_out; Times* d_tm;
    cudaMalloc(&d_in, sizeof(h_in)); cudaMalloc(&d_out, THREADS * 4); cudaMalloc(&d_tm, nsm * sizeof(Times));
    cudaMemcpy(d_in, h_in, sizeof(h_in), cudaMemcpyHostToDevice);
    const double per = (double)CHAINS * UNROLL;
    std::vector<Test> tests = {
        {"popc", t_popc, per, "POPC"},
        {"lop3", t_lop3, per, "LOP3.LUT"},
        {"iadd3", t_iadd3, per, "IADD3"},
        {"imad", t_imad, per, "IMAD (32-bit multiply-add)"},
        {"shf", t_shf, per, "SHF (funnel shift)"},
        {"prmt", t_prmt, per, "PRMT (byte permute)"},
        {"vimnmx3_s16x2", t_vimnmx3, per, "VIMNMX3.S16x2 (3-input packed 16-bit max; FAST arc score)"},
        {"vabsdiff4", t_vabsdiff4, per, "VABSDIFF4.U8 (4 byte differences; FAST necessary test)"},
        {"dp4a", t_dp4a, per, "IDP.4A (4 x u8 dot product; blur rows)"},
        {"dp2a", t_dp2a, per, "IDP.2A (2 x u16 . u8 dot product; resize rows)"},
        {"mix_xor_popc", t_xor_popc, 2 * per, "LOP3 + POPC alternating (counts both)"},
        {"mix_lop3_imad", t_lop3_imad, 2 * per, "LOP3 + IMAD alternating (counts both)"},
        {"mix_dp4a_lop3", t_dp4a_lop3, 2 * per, "IDP.4A + LOP3 alternating (counts both)"},
        {"mix_dp4a_imad", t_dp4a_imad, 2 * per, "IDP.4A + IMAD alternating (counts both)"},
        {"mix_prmt_dp4a", t_prmt_dp4a, 2 * per, "PRMT + IDP.4A alternating (counts both)"},
        {"mix_vimnmx3_lop3", t_vimnmx3_lop3, 2 * per, "VIMNMX3 + LOP3 alternating (counts both)"},
        {"mix_vimnmx3_imad", t_vimnmx3_imad, 2 * per, "VIMNMX3 + IMAD alternating (counts both)"},
        {"hmnmx2", t_hmnmx2, per, "HMNMX2 (half2 max)"},
        {"mix_vimnmx3_hmnmx2", t_vimnmx3_hmnmx2, 2 * per, "VIMNMX3 + HMNMX2 alternating (counts both)"},
        {"mix_hmnmx2_imad", t_hmnmx2_imad, 2 * per, "HMNMX2 + IMAD alternating (counts both)"},
        {"vhmnmx", t_vhmnmx, per, "VHMNMX (three-input half2 max, fused by ptxas from two max.f16x2; one operation each)"},
        {"mix_hmnmx2_lop3", t_hmnmx2_lop3, 2 * per, "HMNMX2 + LOP3 alternating (counts both)"},
        {"hamming256_popc", t_hamming256, 16.0 * 8.0, "256-bit Hamming distance as the matcher does it: 2 LDS.128 + 8 x (LOP3 + POPC + IADD); counts POPCs"},
        {"lds32", t_lds<0>, per, "LDS.32, consecutive words (conflict-free); counts 4-byte lanes"},
        {"lds32_broadcast", t_lds<1>, per, "LDS.32, one word per warp (broadcast)"},
        {"lds_u8", t_lds<2>, per, "LDS.U8, one byte per lane in distinct banks"},
    };
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::vector<Times> tm(nsm);
    const int iters = argc > 1 ? atoi(argv[1]) : 4000;
    printf("{\n \"gpu_name\": \"%s\", \"sms\": %d, \"clock_rate_attr_mhz\": %.1f,\n", pr.name, nsm, clk_khz / 1e3);
    printf(" \"how\": \"tools/peaks_int.cu: one CTA of %d threads per SM, %d independent chains x %d unrolled ops per loop iteration, %d iterations; "
           "lanes_per_clk_per_sm from clock64() inside the kernel (median over SMs), gops from CUDA events; best of 5 runs\",\n", THREADS, CHAINS, UNROLL, iters);
    printf(" \"tests\": {\n");
    double clk_sum = 0; int clk_n = 0;
    for (size_t t = 0; t < tests.size(); ++t) {
        const Test& T = tests[t];
        double best_lanes = 0, best_gops = 0, best_mhz = 0;
        for (int rep = 0; rep < 6; ++rep) {
            cudaEventRecord(e0);
            T.k<<<nsm, THREADS>>>(d_in, d_out, d_tm, iters);
            cudaEventRecord(e1);
            if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "kernel %s failed: %s\n", T.name, cudaGetErrorString(cudaGetLastError())); return 1; }
            if (rep == 0) continue;   // warm-up
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            cudaMemcpy(tm.data(), d_tm, nsm * sizeof(Times), cudaMemcpyDeviceToHost);
            std::vector<double> cyc(nsm);
            for (int i = 0; i < nsm; ++i) cyc[i] = (double)(tm[i].t1 - tm[i].t0);
            std::sort(cyc.begin(), cyc.end());
            const double med = cyc[nsm / 2];
            const double ops_sm = T.ops_per_thread_iter * iters * THREADS;
            const double lanes = ops_sm / med;
            const double gops = ops_sm * nsm / (ms * 1e-3) / 1e9;
            if (lanes > best_lanes) { best_lanes = lanes; best_mhz = med / (ms * 1e-3) / 1e6; }
            best_gops = std::max(best_gops, gops);
        }
        clk_sum += best_mhz; ++clk_n;
        printf("  \"%s\": {\"lanes_per_clk_per_sm\": %.2f, \"gops\": %.1f, \"sm_mhz_implied\": %.0f, \"what\": \"%s\"}%s\n", T.name, best_lanes, best_gops,
               best_mhz, T.what, t + 1 < tests.size() ? "," : "");
    }
    printf(" },\n \"sm_mhz_implied_mean\": %.0f\n}\n", clk_sum / clk_n);
    return 0;
}
