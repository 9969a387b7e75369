// Stand-alone probe of the TMA path used by k_stage1_tma: one 3-D FLOAT64 box load per variant, checked element by element.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <vector>
typedef CUresult (*PFN)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
struct Maps { CUtensorMap m[2]; };
__device__ __forceinline__ unsigned s32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ Maps M, int which, int rows, int cols, int c0, int c1, int c2, double* out, int dst_off) {
    extern __shared__ __align__(128) unsigned char sm[];
    unsigned bar = s32(sm + 32768);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(rows * cols * 8) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"(s32(sm + dst_off)), "l"(reinterpret_cast<unsigned long long>(&M.m[which])), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
    }
    __syncwarp();
    unsigned ok;
    do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(0) : "memory"); } while (!ok);
    const double* s = reinterpret_cast<const double*>(sm + dst_off);
    for (int q = threadIdx.x; q < rows * cols; q += 32) out[q] = s[q];
}
int main() {
    const int nx = 64, ny = 128, ne = 9; const long long N = (long long)nx * ny;
    std::vector<double> h(N * ne); for (size_t q = 0; q < h.size(); q++) h[q] = (double)q;
    double *d, *out; cudaMalloc(&d, h.size() * 8); cudaMalloc(&out, 65536); cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { printf("no entry point\n"); return 1; }
    PFN enc = (PFN)p;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000);
    struct V { int rows, cols, off; CUtensorMapL2promotion l2; const char* name; } vs[] = {
        {32, 8, 0, CU_TENSOR_MAP_L2_PROMOTION_NONE, "32x8 off0 l2none"}, {32, 8, 0, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, "32x8 off0 l2 128"},
        {34, 10, 0, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, "34x10 off0"}, {36, 8, 0, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, "36x8 off0"},
        {32, 10, 2816, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, "32x10 off2816"}, {32, 8, 2720, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, "32x8 off2720 (16B aligned only)"}};
    for (auto& v : vs) {
        Maps M; memset(&M, 0, sizeof(M));
        const cuuint64_t dims[3] = {(cuuint64_t)ny, (cuuint64_t)nx, (cuuint64_t)ne}; const cuuint64_t st[2] = {(cuuint64_t)ny * 8, (cuuint64_t)N * 8};
        const cuuint32_t box[3] = {(cuuint32_t)v.rows, (cuuint32_t)v.cols, 1}, es[3] = {1, 1, 1};
        CUresult r = enc(&M.m[1], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, v.l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("%-36s encode failed %d\n", v.name, (int)r); continue; }
        cudaMemset(out, 0, 65536);
        const int c0 = 31, c1 = 7, c2 = 2;
        probe<<<1, 32, 40000>>>(M, 1, v.rows, v.cols, c0, c1, c2, out, v.off);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%-36s kernel: %s\n", v.name, cudaGetErrorString(e)); return 2; }
        std::vector<double> o(v.rows * v.cols); cudaMemcpy(o.data(), out, o.size() * 8, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int c = 0; c < v.cols; c++) for (int rr = 0; rr < v.rows; rr++) { const double want = (double)(c2 * N + (long long)(c1 + c) * ny + c0 + rr); if (o[c * v.rows + rr] != want) bad++; }
        printf("%-36s ok, %d wrong of %d\n", v.name, bad, v.rows * v.cols);
    }
    return 0;
}
