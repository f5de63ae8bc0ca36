// Probe of cp.async.bulk.tensor.2d ... tile::gather4 on sm_100a (run on the GPU box: nvcc + execute; scripts/gpu_r2t_gather4.sh).
// Question: can a warp fetch 128 parent rows (64-byte rows of a [rows][16] float table, 4 per lane) with ONE instruction, with
// out-of-range row coordinates zero-filled and still counted in the mbarrier's transaction bytes?
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(const CUtensorMap *map, const int *idx, float *out, int row_off) {
    __shared__ __align__(128) float dst[32 * 4 * 16];
    __shared__ uint64_t bar;
    const int lane = threadIdx.x;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&bar)), "r"(32 * 256) : "memory");
    __syncwarp();
    const int r0 = idx[lane * 4 + 0] + row_off, r1 = idx[lane * 4 + 1] + row_off, r2 = idx[lane * 4 + 2] + row_off, r3 = idx[lane * 4 + 3] + row_off;
    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(
                         smem_addr(dst + lane * 64)),
                 "l"(map), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_addr(&bar))
                 : "memory");
    uint32_t done = 0;
    for (int spin = 0; spin < 2000000 && !done; spin++)
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(smem_addr(&bar)) : "memory");
    if (lane == 0 && !done) printf("TIMEOUT: the transaction bytes never completed\n");
    __syncwarp();
    for (int i = lane; i < 32 * 64; i += 32) out[i] = dst[i];
}

int main() {
    const int N = 1024;
    std::vector<float> rows((size_t)N * 16);
    for (int r = 0; r < N; r++)
        for (int c = 0; c < 16; c++) rows[(size_t)r * 16 + c] = (float)(r * 100 + c);
    float *d_rows, *d_out;
    int *d_idx;
    CUtensorMap *d_map;
    cudaMalloc(&d_rows, rows.size() * 4), cudaMalloc(&d_out, 32 * 64 * 4), cudaMalloc(&d_idx, 128 * 4), cudaMalloc(&d_map, sizeof(CUtensorMap));
    cudaMemcpy(d_rows, rows.data(), rows.size() * 4, cudaMemcpyHostToDevice);
    std::vector<int> idx(128);
    for (int i = 0; i < 128; i++) idx[i] = (i * 37 + 11) % N;
    idx[5] = -1, idx[6] = N, idx[7] = N + 12345, idx[64] = 0x7FFFFFF0; // out of range: expect zeros
    cudaMemcpy(d_idx, idx.data(), 128 * 4, cudaMemcpyHostToDevice);

    typedef CUresult (*Encode)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                               const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    CUtensorMap map;
    const cuuint64_t dims[2] = {16, (cuuint64_t)N}, strides[1] = {64};
    const cuuint32_t box[2] = {16, 1}, estr[2] = {1, 1};
    CUresult r = ((Encode)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d_rows, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    if (r != CUDA_SUCCESS) return 1;
    cudaMemcpy(d_map, &map, sizeof map, cudaMemcpyHostToDevice);
    for (int row_off : {0}) {
        probe<<<1, 32>>>(d_map, d_idx, d_out, row_off);
        cudaError_t e = cudaDeviceSynchronize();
        printf("kernel: %s\n", cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        std::vector<float> out(32 * 64);
        cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int i = 0; i < 128; i++) {
            const int rr = idx[i];
            for (int c = 0; c < 16; c++) {
                const float want = (rr >= 0 && rr < N) ? (float)(rr * 100 + c) : 0.0f, got = out[(size_t)i * 16 + c];
                if (want != got && bad++ < 8) printf("row %d (index %d) col %d: got %g want %g\n", i, rr, c, got, want);
            }
        }
        printf("gather4 probe: %d mismatches of %d\n", bad, 128 * 16);
    }
    return 0;
}
