// Probe: how does tcgen05.mma kind::tf32 convert fp32 operands (truncate or round)?
// One 128x256x8 MMA on hand-laid SWIZZLE_NONE K-major tiles; prints D[0][0..3].
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/tf32_probe tools/tf32_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ uint64_t desc(uint32_t saddr) {  // SWIZZLE_NONE, LBO = 128 B (K-adjacent core), SBO = 256 B (next 8 rows)
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(128 >> 4) << 16;
    d |= (uint64_t)(256 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
__global__ void probe(const float* a_vals, const float* b_vals, float* out) {
    __shared__ __align__(1024) float A[128 * 8];
    __shared__ __align__(1024) float B[256 * 8];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tptr;
    // element (row r, k): core group r/8, row-in-core r%8, k-half k/4, k%4
    auto idx = [](int r, int k) { return (r / 8) * 64 + (k / 4) * 32 + (r % 8) * 4 + (k % 4); };
    for (int i = threadIdx.x; i < 128 * 8; i += blockDim.x) A[i] = 0.f;
    for (int i = threadIdx.x; i < 256 * 8; i += blockDim.x) B[i] = 0.f;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 0; k < 8; ++k) A[idx(0, k)] = a_vals[k];
        for (int j = 0; j < 4; ++j)
            for (int k = 0; k < 8; ++k) B[idx(j, k)] = b_vals[j * 8 + k];
        uint32_t b32 = (uint32_t)__cvta_generic_to_shared(&bar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b32));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t t32 = (uint32_t)__cvta_generic_to_shared(&tptr);
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(t32) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem = tptr;
    uint32_t b32 = (uint32_t)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) {
        uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
        uint64_t ad = desc((uint32_t)__cvta_generic_to_shared(A)), bd = desc((uint32_t)__cvta_generic_to_shared(B));
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}"
                     ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(0u) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(b32) : "memory");
    }
    if (threadIdx.x < 32) {
        asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}" ::"r"(b32) : "memory");
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t r0, r1, r2, r3;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(tmem));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (threadIdx.x == 0) {
            out[0] = __uint_as_float(r0); out[1] = __uint_as_float(r1);
            out[2] = __uint_as_float(r2); out[3] = __uint_as_float(r3);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}
int main() {
    float ha[8] = {0}, hb[32] = {0}, ho[4];
    const float u = 1.0f / 1024.0f;          // 2^-10: one TF32 ulp at 1.0
    ha[0] = 1.0f;
    hb[0 * 8 + 0] = 1.0f + 0.75f * u;        // col 0: A exact, B between two TF32 values (3/4 up)
    hb[1 * 8 + 0] = 1.0f + 0.25f * u;        // col 1: B 1/4 up
    hb[2 * 8 + 0] = -(1.0f + 0.75f * u);     // col 2: negative, 3/4
    hb[3 * 8 + 0] = 1.0f + 0.5f * u;         // col 3: exact tie
    float *da, *db, *dout;
    cudaMalloc(&da, sizeof(ha)); cudaMalloc(&db, sizeof(hb)); cudaMalloc(&dout, sizeof(ho));
    cudaMemcpy(da, ha, sizeof(ha), cudaMemcpyHostToDevice); cudaMemcpy(db, hb, sizeof(hb), cudaMemcpyHostToDevice);
    probe<<<1, 128>>>(da, db, dout);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(ho, dout, sizeof(ho), cudaMemcpyDeviceToHost);
    printf("err=%s\n", cudaGetErrorString(e));
    const char* nm[4] = {"B=1+0.75u", "B=1+0.25u", "B=-(1+0.75u)", "B=1+0.5u"};
    for (int j = 0; j < 4; ++j) printf("%-14s D=%.10f  (D-1)/u=%.4f\n", nm[j], ho[j], (fabsf(ho[j]) - 1.0f) / u);
    // now A inexact, B exact
    ha[0] = 1.0f + 0.75f * u; for (int j = 0; j < 4; ++j) hb[j * 8] = 1.0f;
    ha[1] = 3.0f * u * u; hb[0 * 8 + 1] = 1.0f;   // col 0 also gets a tiny second product: accumulation precision
    cudaMemcpy(da, ha, sizeof(ha), cudaMemcpyHostToDevice); cudaMemcpy(db, hb, sizeof(hb), cudaMemcpyHostToDevice);
    probe<<<1, 128>>>(da, db, dout);
    cudaDeviceSynchronize();
    cudaMemcpy(ho, dout, sizeof(ho), cudaMemcpyDeviceToHost);
    printf("A=1+0.75u, B=1: D=%.10f (D-1)/u=%.4f ; with +3u^2 term: D=%.10f (D-1)/u=%.6f\n", ho[1], (ho[1] - 1.0f) / u, ho[0], (ho[0] - 1.0f) / u);
    return 0;
}
