// Development aid: write bandwidth when a 62.7 MB buffer is written as chunks of S bytes whose order
// over the grid is either sequential or "tile-like" (64 rows of a column block, row pitch 6272 B).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void k_seq(uint4* out, size_t n16) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i < n16; i += (size_t)gridDim.x * blockDim.x) out[i] = make_uint4(1, 2, 3, 4);
}
// tile-like: tile t -> (row tile rt, col block cb); warp w writes rows w, w+8, ...; lane 16 B
__global__ void k_tile(uint4* out, int n_rows, int pitch16, int cb_count, int seg16, int tile_rows) {
    int n_tiles = (n_rows / tile_rows) * cb_count;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        int rt = t / cb_count, cb = t % cb_count;
        for (int r = warp; r < tile_rows; r += nw)
            for (int c = lane; c < seg16; c += 32) {
                int col = cb * seg16 + c;
                if (col < pitch16) out[(size_t)(rt * tile_rows + r) * pitch16 + col] = make_uint4(1, 2, 3, 4);
            }
    }
}
template <int MODE>  // 0 = st.global.cs from registers, 1 = smem tile + TMA bulk store, 2 = as 1 but wait before reuse only
__global__ void k_tile2(uint4* out, int n_rows, int pitch16, int cb_count, int seg16, int tile_rows) {
    extern __shared__ __align__(128) uint4 sbuf[];
    int n_tiles = (n_rows / tile_rows) * cb_count;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        int rt = t / cb_count, cb = t % cb_count;
        if (MODE == 0) {
            for (int r = warp; r < tile_rows; r += nw)
                for (int c = lane; c < seg16; c += 32) {
                    int col = cb * seg16 + c;
                    if (col < pitch16) __stcs(&out[(size_t)(rt * tile_rows + r) * pitch16 + col], make_uint4(1, 2, 3, 4));
                }
        } else {
            if (warp == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncthreads();
            for (int r = warp; r < tile_rows; r += nw)
                for (int c = lane; c < seg16; c += 32) sbuf[r * (seg16 + 1) + c] = make_uint4(1, 2, 3, 4);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (warp == 0) {
                int segb = min(seg16, pitch16 - cb * seg16) * 16;
                for (int r = lane; r < tile_rows; r += 32) {
                    void* g = &out[(size_t)(rt * tile_rows + r) * pitch16 + cb * seg16];
                    unsigned sa = (unsigned)__cvta_generic_to_shared(&sbuf[r * (seg16 + 1)]);
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(sa), "r"(segb) : "memory");
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    if (MODE != 0 && warp == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// MODE bits: 1 = __syncthreads twice per tile, 2 = dynamic atomic tile scheduler, 4 = per-tile global loads
// feeding the stored value, 8 = row offsets read from shared memory, 16 = 100 dependent ALU ops per tile
template <int MODE>
__global__ void k_tile3(uint4* out, int n_rows, int pitch16, int cb_count, int seg16, int tile_rows,
                        unsigned* sched, const unsigned* __restrict__ aux, int cols16 = 0) {
    if (cols16 == 0) cols16 = pitch16;
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_off[64];
    __shared__ unsigned s_big[(MODE & 128) ? 9000 : 1];
    if (MODE & 128) s_big[threadIdx.x] = threadIdx.x;
    if (MODE & 64) {
        if (!(MODE & 256)) asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    }
    int n_tiles = (n_rows / tile_rows) * cb_count;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    int t = blockIdx.x;
    for (;;) {
        if (MODE & 2) {
            if (threadIdx.x == 0) s_tile = atomicAdd(&sched[0], 1u);
            __syncthreads();
            t = s_tile;
        }
        if (t >= n_tiles) break;
        int tt = (MODE & 32) ? n_tiles - 1 - t : t;
        int rt = tt / cb_count, cb = tt % cb_count;
        unsigned x = 1;
        if (MODE & 4) x = aux[(t * 131 + threadIdx.x) & 0xffff];
        if (MODE & 16) {
#pragma unroll 1
            for (int i = 0; i < 100; ++i) x = x * 1664525u + 1013904223u;
        }
        if (MODE & 8) {
            if (threadIdx.x < tile_rows) s_off[threadIdx.x] = (unsigned long long)(rt * tile_rows + threadIdx.x) * pitch16;
        }
        if (MODE & 1) __syncthreads();
        for (int r = warp; r < tile_rows; r += nw)
            for (int c = lane; c < seg16; c += 32) {
                int col = cb * seg16 + c;
                size_t off = (MODE & 8) ? s_off[r] : (size_t)(rt * tile_rows + r) * pitch16;
                if (col < cols16) __stcs(&out[off + col], make_uint4(x, 2, 3, 4));
            }
        if (MODE & 1) __syncthreads();
        if (!(MODE & 2)) { t += gridDim.x; }
    }
    if ((MODE & 2) && threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&sched[1], 1u) == gridDim.x - 1) { sched[0] = 0; sched[1] = 0; }
    }
    if (MODE & 256) asm volatile("griddepcontrol.wait;" ::: "memory");  // complete only after the predecessor
}
// a small kernel between two big ones, as k_prep sits between two k_expand launches
template <int WAIT_AT_END>
__global__ void k_mid(unsigned* ctr, int spin_ns) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (threadIdx.x == 0) {
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
        do { asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1)); } while (t1 - t0 < (unsigned long long)spin_ns);
        atomicAdd(ctr, 1u);
        if (WAIT_AT_END) asm volatile("griddepcontrol.wait;" ::: "memory");
    }
}
int main() {
    const int rows = 9984, pitch = 6272;  // bytes
    size_t bytes = (size_t)rows * pitch;
    uint4* buf[4];
    for (auto& b : buf) cudaMalloc(&b, bytes);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto run = [&](const char* name, auto launch) {
        for (int i = 0; i < 5; ++i) launch(buf[i % 4]);
        cudaEventRecord(e0);
        for (int i = 0; i < 50; ++i) launch(buf[i % 4]);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("%-40s %7.1f us  %7.1f GB/s  (%s)\n", name, ms / 50 * 1e3, bytes / (ms / 50 * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
    };
    run("sequential grid-stride 148x8x256", [&](uint4* o) { k_seq<<<148 * 8, 256>>>(o, bytes / 16); });
    run("sequential one-shot", [&](uint4* o) { k_seq<<<(unsigned)(bytes / 16 / 256), 256>>>(o, bytes / 16); });
    int segs[] = {8, 32, 98, 392};  // 16-byte units: 128 B, 512 B, 1568 B, 6272 B
    int trs[] = {64, 16, 8};
    for (int tr : trs)
        for (int s : segs) {
            int cbc = (392 + s - 1) / s;
            char name[96];
            snprintf(name, sizeof name, "tile %d rows x %d B, grid 740", tr, s * 16);
            run(name, [&](uint4* o) { k_tile<<<740, 256>>>(o, rows, 392, cbc, s, tr); });
        }
    for (int grid : {740, 1480}) {
        char name[96];
        snprintf(name, sizeof name, "st.cs regs 64 x 512 B, grid %d", grid);
        run(name, [&](uint4* o) { k_tile2<0><<<grid, 256>>>(o, rows, 392, 13, 32, 64); });
        snprintf(name, sizeof name, "smem+TMA bulk 64 x 512 B, grid %d", grid);
        cudaFuncSetAttribute(k_tile2<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 33 * 16);
        run(name, [&](uint4* o) { k_tile2<1><<<grid, 256, 64 * 33 * 16>>>(o, rows, 392, 13, 32, 64); });
    }
    unsigned* sched; cudaMalloc(&sched, 8); cudaMemset(sched, 0, 8);
    unsigned* aux; cudaMalloc(&aux, 65536 * 4); cudaMemset(aux, 1, 65536 * 4);
#define RUN3(M) run("k_tile3 mode " #M, [&](uint4* o) { k_tile3<M><<<740, 256>>>(o, rows, 392, 13, 32, 64, sched, aux); });
    RUN3(0) RUN3(1) RUN3(9) RUN3(11)
    // the real layout: 10000 rows (156 full row tiles here), pitch 6400 B, 1564 words used
    for (auto& b : buf) { cudaFree(b); cudaMalloc(&b, (size_t)10048 * 6400); }
#define RUN4(M) run("pitch 6400: k_tile3 mode " #M, [&](uint4* o) { k_tile3<M><<<740, 256>>>(o, rows, 400, 13, 32, 64, sched, aux, 391); });
    RUN4(0) RUN4(32) RUN4(9) RUN4(2) RUN4(11) RUN4(43)
    unsigned* sched2; cudaMalloc(&sched2, 64); cudaMemset(sched2, 0, 64);
    int flip = 0;
    int grid2 = 740;
    auto pdl2 = [&](auto kern, uint4* o) {  // consecutive launches overlap: each has its own ticket counter
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid2); cfg.blockDim = dim3(256);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        flip = (flip + 1) % 3;
        cudaLaunchKernelEx(&cfg, kern, o, rows, 400, 13, 32, 64, sched2 + 4 * flip, (const unsigned*)aux, 392);
    };
    auto pdl = [&](auto kern, uint4* o) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(740); cfg.blockDim = dim3(256);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, kern, o, rows, 400, 13, 32, 64, sched, (const unsigned*)aux, 392);
    };
    run("pitch 6400/392: static", [&](uint4* o) { k_tile3<0><<<740, 256>>>(o, rows, 400, 13, 32, 64, sched, aux, 392); });
    run("pitch 6400/392: static + 36KB smem", [&](uint4* o) { k_tile3<128><<<740, 256>>>(o, rows, 400, 13, 32, 64, sched, aux, 392); });
    run("pitch 6400/392: static + PDL", [&](uint4* o) { pdl(k_tile3<64>, o); });
    run("pitch 6400/392: static + PDL + smem", [&](uint4* o) { pdl(k_tile3<64 + 128>, o); });
    run("pitch 6400/392: tickets", [&](uint4* o) { k_tile3<2><<<740, 256>>>(o, rows, 400, 13, 32, 64, sched, aux, 392); });
    run("pitch 6400/392: tickets + PDL + smem", [&](uint4* o) { pdl(k_tile3<2 + 64 + 128>, o); });
    run("pitch 6400/392: tickets + PDL overlapped", [&](uint4* o) { pdl2(k_tile3<2 + 64 + 256>, o); });
    run("pitch 6400/392: tickets + PDL overlapped + smem", [&](uint4* o) { pdl2(k_tile3<2 + 64 + 128 + 256>, o); });
    unsigned* ctr; cudaMalloc(&ctr, 4); cudaMemset(ctr, 0, 4);
    auto mid = [&](auto kern, int blocks) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(128);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, kern, ctr, 2000);
    };
    run("overlapped + mid kernel 148 blocks", [&](uint4* o) { mid(k_mid<0>, 148); pdl2(k_tile3<2 + 64 + 128 + 256>, o); });
    run("overlapped + mid kernel 470 blocks", [&](uint4* o) { mid(k_mid<0>, 470); pdl2(k_tile3<2 + 64 + 128 + 256>, o); });
    run("overlapped + mid kernel 470, wait at end", [&](uint4* o) { mid(k_mid<1>, 470); pdl2(k_tile3<2 + 64 + 128 + 256>, o); });
    for (int g : {592, 444, 296}) {
        grid2 = g;
        char name[96];
        snprintf(name, sizeof name, "overlapped, grid %d", g);
        run(name, [&](uint4* o) { pdl2(k_tile3<2 + 64 + 128 + 256>, o); });
    }
    grid2 = 740;
    // row pitch sweep, 391 x 16 B used per row, static round-robin tiles of 64 rows x 512 B
    for (auto& b : buf) { cudaFree(b); cudaMalloc(&b, (size_t)10048 * 8192); }
    for (int p16 : {391, 392, 394, 396, 400, 404, 408, 416, 424, 448, 512}) {
        char name[96];
        snprintf(name, sizeof name, "pitch %d B static", p16 * 16);
        run(name, [&](uint4* o) { k_tile3<0><<<740, 256>>>(o, rows, p16, 13, 32, 64, sched, aux, 391); });
    }
    return 0;
}
