// Development aid: what a "one warp = one tile, registers only" bitmap writer can reach on the C4 shape.
// Each warp owns one tile of 64 rows x 512 B (row pitch 6272 B, 13 column blocks, 157 row tiles): it loads what
// the real kernel would load for the tile (ten 512-byte bucket planes, the 64 rows' offsets), combines them with a
// few logic ops and stores the tile's rows from registers; no shared memory, no block-level barrier, no tickets.
// Variants: warps per CTA, loads on/off, row permutation, consecutive launches overlapped (programmatic dependent
// launch without a wait at the start).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/warp_tile_bw.cu -o tools/warp_tile_bw
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include <type_traits>

constexpr int kRows = 10000, kPitchW = 1568, kWordsOut = 1568, kCB = 13;

template <int MODE, int kTileRows>  // bit 0: tile loads; bit 1: PDL prologue; bit 2: complete only after the predecessor
__global__ void k_warp_tile(uint32_t* __restrict__ out, const uint32_t* __restrict__ planes, const uint16_t* __restrict__ perm,
                            int n_tiles) {
    const int lane = threadIdx.x & 31;
    const int tile_lin = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (MODE & 2) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (tile_lin >= n_tiles) return;
    // consecutive warps (the warps of a CTA) take consecutive row tiles of ONE column block: they share the planes
    const int n_rt = (kRows + kTileRows - 1) / kTileRows;
    const int tile = n_tiles - 1 - tile_lin;
    const int cb = tile / n_rt, rt = tile % n_rt;
    const int j0 = rt * kTileRows;
    const int rows = min(kTileRows, kRows - j0);
    uint4 b4 = make_uint4(1, 2, 3, 4);
    uint32_t row_a = min(j0 + lane, kRows - 1), row_b = min(j0 + 32 + lane, kRows - 1);
    if (MODE & 1) {
        uint4 acc = make_uint4(~0u, ~0u, ~0u, ~0u);
#pragma unroll
        for (int k = 0; k < 10; ++k) {
            const uint4 p = *reinterpret_cast<const uint4*>(planes + ((size_t)cb * 10 + k) * 128 + lane * 4);
            if ((rt >> k) & 1) { acc.x &= p.x; acc.y &= p.y; acc.z &= p.z; acc.w &= p.w; }
            else { acc.x |= p.x; acc.y |= p.y; acc.z |= p.z; acc.w |= p.w; }
        }
        b4 = acc;
        row_a = perm[min(j0 + lane, kRows - 1)];
        row_b = perm[min(j0 + 32 + lane, kRows - 1)];
    }
    const int col = cb * 128 + lane * 4;
    if (col < kWordsOut) {
        uint32_t* dst = out + col;
#pragma unroll 8
        for (int r = 0; r < rows; ++r) {
            const uint32_t row = __shfl_sync(0xffffffffu, r < 32 ? row_a : row_b, r & 31);
            __stcs(reinterpret_cast<uint4*>(dst + (size_t)row * kPitchW), b4);
        }
    } else {
        for (int r = 0; r < rows; ++r) (void)__shfl_sync(0xffffffffu, r < 32 ? row_a : row_b, r & 31);
    }
    if (MODE & 4) asm volatile("griddepcontrol.wait;" ::: "memory");
}

int main() {
    const size_t bytes = (size_t)kRows * kPitchW * 4;
    uint32_t* buf[4];
    for (auto& b : buf) cudaMalloc(&b, bytes);
    uint32_t* planes;
    cudaMalloc(&planes, kCB * 10 * 128 * 4);
    cudaMemset(planes, 0x5a, kCB * 10 * 128 * 4);
    uint16_t* perm;
    cudaMalloc(&perm, kRows * 2);
    {
        uint16_t* h = new uint16_t[kRows];
        for (int i = 0; i < kRows; ++i) h[i] = (uint16_t)((i * 7919) % kRows);  // a permutation (7919 and 10000 coprime)
        cudaMemcpy(perm, h, kRows * 2, cudaMemcpyHostToDevice);
        delete[] h;
    }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    auto run = [&](const char* name, auto launch) {
        for (int i = 0; i < 5; ++i) launch(buf[i % 4]);
        cudaDeviceSynchronize();
        float best = 1e9f, sum = 0;
        for (int rep = 0; rep < 5; ++rep) {
            cudaEventRecord(e0);
            for (int i = 0; i < 50; ++i) launch(buf[i % 4]);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            best = ms < best ? ms : best;
            sum += ms;
        }
        printf("%-56s best %6.2f us  mean %6.2f us  %7.1f GB/s  (%s)\n", name, best / 50 * 1e3, sum / 250 * 1e3,
               bytes / (best / 50 * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
    };
    auto sweep = [&](auto tr_tag) {
        constexpr int TR = decltype(tr_tag)::value;
        const int n_tiles = ((kRows + TR - 1) / TR) * kCB;
        for (int wpc : {1, 4, 8}) {
            const int grid = (n_tiles + wpc - 1) / wpc;
            char name[128];
            snprintf(name, sizeof name, "%2d rows/warp, %d warp(s)/CTA, stores only", TR, wpc);
            run(name, [&](uint32_t* o) { k_warp_tile<0, TR><<<grid, wpc * 32>>>(o, planes, perm, n_tiles); });
            snprintf(name, sizeof name, "%2d rows/warp, %d warp(s)/CTA, loads + permuted rows", TR, wpc);
            run(name, [&](uint32_t* o) { k_warp_tile<1, TR><<<grid, wpc * 32>>>(o, planes, perm, n_tiles); });
            auto pdl = [&](auto kern, uint32_t* o) {
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(grid);
                cfg.blockDim = dim3(wpc * 32);
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                at[0].val.programmaticStreamSerializationAllowed = 1;
                cfg.attrs = at;
                cfg.numAttrs = 1;
                cudaLaunchKernelEx(&cfg, kern, o, (const uint32_t*)planes, (const uint16_t*)perm, n_tiles);
            };
            snprintf(name, sizeof name, "%2d rows/warp, %d warp(s)/CTA, loads, launches overlapped", TR, wpc);
            run(name, [&](uint32_t* o) { pdl(k_warp_tile<1 + 2 + 4, TR>, o); });
        }
    };
    sweep(std::integral_constant<int, 64>{});
    sweep(std::integral_constant<int, 32>{});
    sweep(std::integral_constant<int, 16>{});
    sweep(std::integral_constant<int, 8>{});
    return 0;
}
