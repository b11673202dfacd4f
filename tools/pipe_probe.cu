// Issue-rate probe for the candidate-flag idioms of the assign epilogue (sm_100a).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/pipe_probe tools/pipe_probe.cu
// Prints cycles per (value, warp) for W warps resident on one SM sub-partition set (one CTA on one SM).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int J>
__device__ __forceinline__ void op_setp_imad(float v, float thr, uint32_t one, uint32_t& acc) {
    constexpr uint32_t K = 1u | ((uint32_t)J << 8) | ((uint32_t)(J * J) << 17);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\t@p mad.lo.u32 %0, %3, %4, %0;\n\t}"
                 : "+r"(acc) : "f"(v), "f"(thr), "r"(one), "n"(K));
}
template <int J>
__device__ __forceinline__ void op_setp_iadd(float v, float thr, uint32_t one, uint32_t& acc) {
    constexpr uint32_t K = 1u | ((uint32_t)J << 8) | ((uint32_t)(J * J) << 17);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\t@p add.u32 %0, %0, %3;\n\t}"
                 : "+r"(acc) : "f"(v), "f"(thr), "n"(K));
}
template <int J>
__device__ __forceinline__ void op_ffma_sat(float v, float thrH, float& acc) {
    // f = sat((thr - v) * 2^100) in {0, 1}; acc += f * K  (both FFMA with an immediate operand)
    constexpr float K = (float)(J + 64 * 128 + 0 * J);   // some per-column constant
    float f;
    asm volatile("fma.rn.sat.f32 %0, %1, 0fCD800000, %2;" : "=f"(f) : "f"(v), "f"(thrH));   // -2^28 (probe value)
    asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc) : "f"(f), "f"(K));
}
template <int J>
__device__ __forceinline__ void op_fset_ffma(float v, float thr, float& acc) {
    constexpr float K = (float)(J + 64 * 128);
    float f;
    asm volatile("set.le.f32.f32 %0, %1, %2;" : "=f"(f) : "f"(v), "f"(thr));
    asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc) : "f"(f), "f"(K));
}
__device__ __forceinline__ float fmin3(float a, float b, float c) {
    float m;
    asm volatile("min.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(a), "f"(b), "f"(c));
    return m;
}

template <int MODE, int I>
struct Body {
    __device__ static __forceinline__ void run(const float (&v)[64], float thr, uint32_t one, uint32_t (&ia)[4],
                                               float (&fa)[4]) {
        if (MODE == 0) op_setp_imad<I>(v[I], thr, one, ia[I & 3]);
        if (MODE == 1) op_ffma_sat<I>(v[I], thr, fa[I & 3]);
        if (MODE == 2) op_fset_ffma<I>(v[I], thr, fa[I & 3]);
        if (MODE == 3) op_setp_iadd<I>(v[I], thr, one, ia[I & 3]);
        if (MODE == 5) {   // mixed: one third on setp+imad, two thirds on the FMA pipe
            if (I % 3 == 0) op_setp_imad<I>(v[I], thr, one, ia[I & 3]);
            else op_ffma_sat<I>(v[I], thr, fa[I & 3]);
        }
        Body<MODE, I + 1>::run(v, thr, one, ia, fa);
    }
};
template <int MODE>
struct Body<MODE, 64> {
    __device__ static __forceinline__ void run(const float (&)[64], float, uint32_t, uint32_t (&)[4], float (&)[4]) {}
};

template <int MODE>
__global__ void probe(const float* in, float* out, long long* cyc, int reps, uint32_t one) {
    float v[64];
    for (int i = 0; i < 64; ++i) v[i] = in[(threadIdx.x * 64 + i) & 4095];
    uint32_t ia[4] = {0, 0, 0, 0};
    float fa[4] = {0, 0, 0, 0};
    float thr = in[threadIdx.x & 4095];
    float mnacc = 3.4e38f;
    __syncthreads();
    const long long c0 = clock64();
    for (int r = 0; r < reps; ++r) {
        if (MODE == 4) {
            float t[22];
#pragma unroll
            for (int k = 0; k < 21; ++k) t[k] = fmin3(v[3 * k], v[3 * k + 1], v[3 * k + 2]);
            t[21] = v[63];
            float u0 = fmin3(t[0], t[1], t[2]), u1 = fmin3(t[3], t[4], t[5]), u2 = fmin3(t[6], t[7], t[8]);
            float u3 = fmin3(t[9], t[10], t[11]), u4 = fmin3(t[12], t[13], t[14]);
            float u5 = fmin3(t[15], t[16], t[17]), u6 = fmin3(t[18], t[19], t[20]);
            float w0 = fmin3(u0, u1, u2), w1 = fmin3(u3, u4, u5);
            mnacc = fmin3(w0, w1, fminf(u6, fminf(t[21], mnacc)));
            v[r & 63] += 1.0f;   // keep the tree from being hoisted
        } else {
            Body<MODE, 0>::run(v, thr, one, ia, fa);
            thr += 0.25f;
        }
    }
    const long long c1 = clock64();
    __syncthreads();
    out[blockIdx.x * blockDim.x + threadIdx.x] =
        __uint_as_float(ia[0] + ia[1] + ia[2] + ia[3]) + fa[0] + fa[1] + fa[2] + fa[3] + mnacc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = c1 - c0;
}

template <int MODE>
static void run(const char* name, const float* in, float* out, long long* cyc) {
    const int reps = 2000;
    printf("%-34s", name);
    for (int warps : {4, 8, 16, 32}) {
        probe<MODE><<<1, warps * 32>>>(in, out, cyc, reps, 1u);
        probe<MODE><<<1, warps * 32>>>(in, out, cyc, reps, 1u);
        cudaDeviceSynchronize();
        long long c;
        cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        // cycles per value per sub-partition: a sub-partition hosts warps/4 warps
        const double per_val = (double)c / ((double)reps * 64.0 * (warps / 4));
        printf("  W=%2d: %.2f cyc/value/SMSP", warps, per_val);
    }
    printf("\n");
}

int main() {
    float *in, *out;
    long long* cyc;
    cudaMalloc(&in, 4096 * 4);
    cudaMalloc(&out, 1024 * 4);
    cudaMalloc(&cyc, 64);
    float h[4096];
    for (int i = 0; i < 4096; ++i) h[i] = (float)((i * 2654435761u) >> 8) * 1e-6f;
    cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
    run<0>("setp.le + @p mad.lo (imm)", in, out, cyc);
    run<3>("setp.le + @p add.u32 (imm)", in, out, cyc);
    run<1>("fma.sat (imm) + fma (imm)", in, out, cyc);
    run<2>("set.le.f32 + fma (imm)", in, out, cyc);
    run<5>("1/3 setp+imad, 2/3 fma.sat+fma", in, out, cyc);
    run<4>("min3 tree (22 + 10 per 64)", in, out, cyc);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
