// FAST over groups of horizontally adjacent cells (included by orb_kernels.cuh).
//
// The reference calls cv::FAST once (twice) per 30-px cell on a ROI that overlaps its neighbours by 6 px; the
// detection zones of the cells of one cell row (ROI minus the 3-px frame) tile the row without gaps. Everything that
// does not depend on the cell — the compass pre-test and the arc score of a pixel — is therefore evaluated on the
// union of up to FASTG_CELLS neighbouring zones at once: one TMA box, full-width rows for the byte-SIMD compass pass,
// one candidate list, one set of barriers. What IS per cell stays per cell:
//   * NMS is blind across the seams: a neighbour pixel outside the own cell's zone counts as "no corner", exactly as
//     cv::FAST on the cell's ROI never scores it;
//   * the threshold choice (ini, else min when the cell is empty) and the row-major ranks (cv::FAST output order)
//     come from per-cell bitmaps and their popc prefix, one warp per cell.
// A CTA walks FASTG_GPB groups; the tile of the next group is fetched by TMA while the current one is processed.
#pragma once

namespace snk
{
constexpr int FASTG_CELLS = 4;    // cells per group
constexpr int FASTG_GW    = 160;  // largest ROI width of a group
constexpr int FASTG_PITCH = 176;  // tile row pitch == TMA box width: group ROI + up to 15 bytes of alignment
constexpr int FASTG_NT    = 256;
constexpr int FASTG_GPB   = 4;    // groups per CTA

struct GroupDev  // 48 bytes
{
    short level, ncell;
    short x0, y0;              // ROI origin of the first cell (level coordinates)
    short h, box_h;            // ROI rows, rows of the level's TMA box
    short off_x, off_y;        // group ROI coordinates -> candidate coordinates (relative to (16,16))
    int cell0;                 // first cell of the group in the image's cell list
    short zb[FASTG_CELLS + 1]; // zone boundaries in group ROI coordinates: zone of cell c = [zb[c], zb[c+1])
    unsigned short zn_word0;   // first tile word that intersects the zones
    unsigned short zn_words;   // number of such words
    unsigned int zn_inv;       // (1 << 20) / zn_words + 1
    short zend;                // == zb[ncell]
    short _pad[3];
};
static_assert(sizeof(GroupDev) == 48, "GroupDev layout");

// CHUNK (SNK_FAST_CHUNK=1, an A/B variant): the compass pass takes 16 pixels (one aligned 16-byte chunk of a tile row) per
// task instead of 4 (one word) — 5 loads instead of 20 per 16 pixels, no per-word index arithmetic, one scan and one atomic
// per warp — and the group's zn_word0 / zn_words / zn_inv then count chunks (orb_extract.cu). Parity green, but 1.179 ms
// against 1.137 ms per 512 images: the loop over the set bits of a lane's 16-pixel mask runs as long as the warp's densest
// lane (corners cluster), which costs more than the arithmetic it saves.
template <bool CHUNK>
__global__ void __launch_bounds__(FASTG_NT) k_fast_group(PyrDev P, const __grid_constant__ TmaMaps maps,
                                                         const GroupDev* __restrict__ groups, int groups_per_img,
                                                         uint32_t* __restrict__ cand, int* __restrict__ cell_count,
                                                         int* __restrict__ err, int tile_rows, int list_cap)
{
    // dynamic shared memory: 2 tile buffers + score tile (tile_rows x FASTG_PITCH each), 2 candidate lists
    extern __shared__ __align__(128) unsigned char fast_smem[];
    const int tile_bytes = tile_rows * FASTG_PITCH;
    uint8_t(*score)[FASTG_PITCH] = reinterpret_cast<uint8_t(*)[FASTG_PITCH]>(fast_smem + 2 * tile_bytes);
    uint16_t* list1 = reinterpret_cast<uint16_t*>(fast_smem + 3 * tile_bytes);
    uint16_t* list2 = list1 + list_cap;
    __shared__ uint32_t bm[FASTG_CELLS][2][FAST_BMW];  // per cell: [0] survivors at the lower threshold, [1] above th_ini
    __shared__ uint16_t pre[FASTG_CELLS][FAST_BMW + 1];
    __shared__ int s_n1, s_n2, s_ini[FASTG_CELLS];
    __shared__ __align__(8) unsigned long long tma_bar[2];

    const int img = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int g0 = blockIdx.x * FASTG_GPB, ng = min(FASTG_GPB, groups_per_img - g0);
    const unsigned bar_addr0  = (unsigned)__cvta_generic_to_shared(&tma_bar[0]);
    const unsigned tile_addr0 = (unsigned)__cvta_generic_to_shared(fast_smem);

    // one TMA box copy per group, completion on the buffer's mbarrier; the box starts at the 16-byte aligned column
    // left of the ROI (unaligned innermost coordinates fault on sm_100a, tools/tma_check.cu)
    auto issue = [&](const GroupDev& gg, int buf) {
        const unsigned bar = bar_addr0 + 8u * buf, dst = tile_addr0 + (unsigned)(buf * tile_bytes);
        const unsigned bytes = (unsigned)(FASTG_PITCH * gg.box_h);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
            "l"(reinterpret_cast<unsigned long long>(&maps.level[gg.level])), "r"((int)gg.x0 - (gg.x0 & 15)), "r"((int)gg.y0), "r"(img),
            "r"(bar)
            : "memory");
    };
    if (tid == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr0 + 8u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    GroupDev g = groups[g0];
    if (tid == 0) issue(g, 0);
    for (int i = tid; i < tile_bytes >> 4; i += FASTG_NT) reinterpret_cast<uint4*>(&score[0][0])[i] = make_uint4(0u, 0u, 0u, 0u);

#pragma unroll 1
    for (int kg = 0; kg < ng; ++kg)
    {
    const int buf = kg & 1;
    uint8_t(*tile)[FASTG_PITCH] = reinterpret_cast<uint8_t(*)[FASTG_PITCH]>(fast_smem + buf * tile_bytes);
    GroupDev gnext = g;
    if (kg + 1 < ng)
    {
        gnext = groups[g0 + kg + 1];
        if (tid == 0) issue(gnext, buf ^ 1);  // that buffer's last readers passed the barrier that ended iteration kg-1
    }
    const int a  = g.x0 & 15;
    const int nc = g.ncell;
    const int zh = g.h - 6;                      // zone rows [3, 3 + zh)
    const int zx0 = g.zb[0], zx1 = g.zend;       // union of the zones, group ROI columns
    // while the copy is in flight: reset the per-group state and the bitmaps (the score tile is zero: cleared once
    // before the first group, and every group resets the entries it wrote)
    if (tid == 0) s_n1 = s_n2 = 0;
    if (tid < FASTG_CELLS) s_ini[tid] = 0;
    for (int i = tid; i < FASTG_CELLS * 2 * FAST_BMW; i += FASTG_NT) (&bm[0][0][0])[i] = 0;
    if (wid == 0)
    {
        // bounded wait: a copy that never completes (bad descriptor) must not hang the device; it is reported instead.
        // One warp polls, the others wait at the barrier below: a polling warp burns issue slots, the kernel's scarce
        // resource, and the first tile of a CTA is a full DRAM latency away
        const unsigned bar = bar_addr0 + 8u * buf, parity = (unsigned)((kg >> 1) & 1);
        unsigned done = 0;
        for (int spin = 0; !done && spin < (1 << 22); ++spin)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x989680;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done)
                         : "r"(bar), "r"(parity)
                         : "memory");
        if (!done && tid == 0) atomicExch(err, 4);
    }
    __syncthreads();
    const int t = P.th_min < P.th_ini ? P.th_min : P.th_ini;

    // ---- A: compass test, 4 pixels (one aligned tile word) per task, tasks = zone rows x words of the zone span ----
    // |p - v| per byte with VABSDIFF4; "byte > t" as bit 7 of ((x & 0x7f) + (127 - t)) | x  (no carry between bytes);
    // a corner needs (S or N beyond the threshold) and (E or W beyond the threshold).
    if (CHUNK && zh > 0 && zx1 > zx0)
    {
        // 16 pixels per task: the centre, south and north chunks as three 16-byte loads, the words left and right of the
        // centre chunk for the x-3 / x+3 bytes; four flag nibbles -> a 16-bit mask of the task's candidates; the list
        // positions from one warp scan and ONE atomic per warp, the entries by a loop over the mask's set bits
        const int c0 = a + zx0, c1 = a + zx1;  // zone columns [c0, c1) in tile coordinates
        const int qa = g.zn_word0, nq = g.zn_words, ntask = zh * nq;
        const unsigned addc = (unsigned)(127 - min(t, 127)) * 0x01010101u;
        const unsigned first_mask = 0xffffu << max(c0 - 16 * qa, 0);
        const unsigned last_mask  = 0xffffu >> max(16 * (qa + nq) - c1, 0);
        const int niter = (ntask + FASTG_NT - 1) / FASTG_NT;
        for (int it = 0; it < niter; ++it)
        {
            const int task = it * FASTG_NT + tid;
            unsigned mask16 = 0;
            int e0 = 0;
            if (task < ntask)
            {
                const int ry = (int)(((unsigned)task * g.zn_inv) >> 20);  // task / nq
                const int qr = task - ry * nq;
                const int y  = 3 + ry;
                const uint8_t* rc = &tile[y][16 * (qa + qr)];
                const uint4 V = *reinterpret_cast<const uint4*>(rc);
                const uint4 S = *reinterpret_cast<const uint4*>(rc + 3 * FASTG_PITCH);
                const uint4 N = *reinterpret_cast<const uint4*>(rc - 3 * FASTG_PITCH);
                // unclamped as in the word-wise pass: foreign bytes only reach pixels the zone mask drops
                const unsigned Lw = *reinterpret_cast<const unsigned*>(rc - 4), Rw = *reinterpret_cast<const unsigned*>(rc + 16);
                const unsigned v[6] = {Lw, V.x, V.y, V.z, V.w, Rw}, s4[4] = {S.x, S.y, S.z, S.w}, n4[4] = {N.x, N.y, N.z, N.w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    const unsigned c  = v[j + 1];
                    const unsigned pE = __funnelshift_r(c, v[j + 2], 24);  // bytes x+3
                    const unsigned pW = __funnelshift_r(v[j], c, 8);       // bytes x-3
                    const unsigned dS = __vabsdiffu4(s4[j], c), dN = __vabsdiffu4(n4[j], c), dE = __vabsdiffu4(pE, c), dW = __vabsdiffu4(pW, c);
                    const unsigned fS = ((dS & 0x7f7f7f7fu) + addc) | dS, fN = ((dN & 0x7f7f7f7fu) + addc) | dN;
                    const unsigned fE = ((dE & 0x7f7f7f7fu) + addc) | dE, fW = ((dW & 0x7f7f7f7fu) + addc) | dW;
                    const unsigned m4 = ((fS | fN) & (fE | fW) & 0x80808080u) >> 7;  // bit 8j: pixel j of the word is a candidate
                    mask16 |= ((m4 * 0x01020408u) >> 24) << (4 * j);                 // bits 24..27 of the product: the four flags
                }
                mask16 &= (qr == 0 ? first_mask : 0xffffu) & (qr == nq - 1 ? last_mask : 0xffffu);
                e0 = (y << 8) + (16 * (qa + qr) - a);  // (the chunk may start left of the ROI: a sum, not an OR)
            }
            const int cnt  = __popc(mask16);
            const int incl = warp_incl_scan(cnt, lane);
            const int tot  = __shfl_sync(0xffffffffu, incl, 31);
            if (tot)
            {
                int base = 0;
                if (lane == 31) base = atomicAdd(&s_n1, tot);
                int off = __shfl_sync(0xffffffffu, base, 31) + incl - cnt;
                while (mask16)
                {
                    const int b = __ffs(mask16) - 1;
                    mask16 &= mask16 - 1;
                    list1[off++] = (uint16_t)(e0 + b);
                }
            }
        }
    }
    if (!CHUNK && zh > 0 && zx1 > zx0)
    {
        const int c0 = a + zx0, c1 = a + zx1;  // zone columns [c0, c1) in tile coordinates
        const int wz0 = g.zn_word0, nwz = g.zn_words, ntask = zh * nwz;
        const unsigned addc = (unsigned)(127 - min(t, 127)) * 0x01010101u;
        // bytes of the first / last word of the span that lie inside the zone columns (all bytes of the words between)
        const int wl = wz0 + nwz - 1;
        const unsigned mfirst = (4 * wz0 < c0) ? (0x80808080u << (8 * (c0 - 4 * wz0))) : 0x80808080u;
        const unsigned mlast  = (4 * wl + 4 > c1) ? (0x80808080u >> (8 * (4 * wl + 4 - c1))) : 0x80808080u;
        const uint32_t* tw = reinterpret_cast<const uint32_t*>(&tile[3][0]) + wz0;
        for (int task = tid; task < ntask; task += FASTG_NT)
        {
            const int ry = (int)(((unsigned)task * g.zn_inv) >> 20);  // task / nwz
            const int wr = task - ry * nwz;
            const int wx = wz0 + wr;
            const int y  = 3 + ry;
            const unsigned zmask = (wr == 0 ? mfirst : 0x80808080u) & (wr == nwz - 1 ? mlast : 0x80808080u);
            // the words left / right of the row are read unclamped: they lie inside the tile (rows 3.. of it), and a
            // zone pixel's x-3 / x+3 never leave the ROI, so foreign bytes only reach pixels the zone mask drops
            const uint32_t* rc = tw + ry * (FASTG_PITCH / 4) + wr;
            const unsigned v  = rc[0];
            const unsigned pS = rc[3 * (FASTG_PITCH / 4)], pN = rc[-3 * (FASTG_PITCH / 4)];
            const unsigned pE = __funnelshift_r(v, rc[1], 24);   // bytes x+3
            const unsigned pW = __funnelshift_r(rc[-1], v, 8);   // bytes x-3
            const unsigned dS = __vabsdiffu4(pS, v), dN = __vabsdiffu4(pN, v), dE = __vabsdiffu4(pE, v), dW = __vabsdiffu4(pW, v);
            const unsigned fS = ((dS & 0x7f7f7f7fu) + addc) | dS, fN = ((dN & 0x7f7f7f7fu) + addc) | dN;
            const unsigned fE = ((dE & 0x7f7f7f7fu) + addc) | dE, fW = ((dW & 0x7f7f7f7fu) + addc) | dW;
            unsigned m4 = (fS | fN) & (fE | fW) & zmask;  // bit 8j+7: pixel j of the word is a candidate
            if (m4)
            {
                int off = atomicAdd(&s_n1, __popc(m4));  // list order is irrelevant (ranks come from the bitmaps)
                const int e0 = (y << 8) | (4 * wx - a);  // four predicated stores beat a loop over the set bits
                if (m4 & 0x00000080u) list1[off++] = (uint16_t)e0;
                if (m4 & 0x00008000u) list1[off++] = (uint16_t)(e0 + 1);
                if (m4 & 0x00800000u) list1[off++] = (uint16_t)(e0 + 2);
                if (m4 & 0x80000000u) list1[off] = (uint16_t)(e0 + 3);
            }
        }
    }
    __syncthreads();
    // ---- B: arc score of every candidate; corners (arc > lower threshold) go to list2 ----
    const int n1 = s_n1;
    for (int base = 0; base < n1; base += FASTG_NT)
    {
        const int k = base + tid;
        bool pass   = false;
        uint16_t e  = 0;
        if (k < n1)
        {
            e = list1[k];
            const int y = e >> 8, xa = (e & 255) + a;
            const int arc = fast_arc_score<FASTG_PITCH>(tile, y, xa);
            pass = arc > t;
            if (pass) score[y][xa] = (uint8_t)arc;
        }
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        int off = 0;
        if (lane == 0 && m) off = atomicAdd(&s_n2, __popc(m));
        off = __shfl_sync(0xffffffffu, off, 0);
        if (pass) list2[off + __popc(m & ((1u << lane) - 1))] = e;
    }
    __syncthreads();
    const int n2 = s_n2;
    // cell of a group ROI column and its zone [zs, ze): unused boundaries hold 32767
    const int b1 = g.zb[1], b2 = g.zb[2], b3 = g.zb[3], b4 = g.zb[4];
    auto cell_of = [&](int x, int& zs, int& ze) {
        const int c = (x >= b1) + (x >= b2) + (x >= b3);
        zs = (c == 0) ? zx0 : (c == 1) ? b1 : (c == 2) ? b2 : b3;
        ze = (c == 0) ? b1 : (c == 1) ? b2 : (c == 2) ? b3 : b4;
        ze = min(ze, zx1);
        return c;
    };
    // ---- C: NMS inside the own cell's zone, survivors into the cell's bitmaps ----
    for (int k = tid; k < n2; k += FASTG_NT)
    {
        const int y = list2[k] >> 8, x = list2[k] & 255, xa = x + a;
        int zs, ze;
        const int c = cell_of(x, zs, ze);
        const int s = score[y][xa];
        // neighbours beyond the cell's zone columns were never scored by the cell's cv::FAST call: they count as 0
        const int mL = (x - 1 >= zs) ? 0xff : 0, mR = (x + 1 < ze) ? 0xff : 0;
        // cv response is s-1 and must beat 0 for empty neighbours: s >= 2
        if (s > 1 && s > (score[y - 1][xa - 1] & mL) && s > score[y - 1][xa] && s > (score[y - 1][xa + 1] & mR) &&
            s > (score[y][xa - 1] & mL) && s > (score[y][xa + 1] & mR) && s > (score[y + 1][xa - 1] & mL) &&
            s > score[y + 1][xa] && s > (score[y + 1][xa + 1] & mR))
        {
            const int i = (y - 3) * (ze - zs) + (x - zs);
            atomicOr(&bm[c][0][i >> 5], 1u << (i & 31));
            if (s > P.th_ini)
            {
                atomicOr(&bm[c][1][i >> 5], 1u << (i & 31));
                s_ini[c] = 1;
            }
        }
    }
    __syncthreads();
    // ---- threshold choice per cell ("if (cv_KeysCell.empty()) retry with th_fast_min") and row-major ranks:
    //      warp c takes cell c, exclusive popc prefix over the cell's bitmap words ----
    if (wid < nc)
    {
        int zs, ze;
        {
            const int c = wid;
            zs = (c == 0) ? zx0 : (c == 1) ? b1 : (c == 2) ? b2 : b3;
            ze = min((c == 0) ? b1 : (c == 1) ? b2 : (c == 2) ? b3 : b4, zx1);
        }
        const int nz = max(ze - zs, 0) * max(zh, 0), nwords = (nz + 31) >> 5;  // <= FAST_BMW
        const uint32_t* B = bm[wid][s_ini[wid] ? 1 : 0];
        int running = 0;
        for (int w0 = 0; w0 < nwords; w0 += 32)
        {
            const int w   = w0 + lane;
            const int cnt = (w < nwords) ? __popc(B[w]) : 0;
            const int inc = warp_incl_scan(cnt, lane);
            if (w < nwords) pre[wid][w] = (uint16_t)(running + inc - cnt);
            running += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0)
        {
            pre[wid][FAST_BMW] = (uint16_t)running;  // total
            cell_count[(size_t)img * P.cells_per_img + g.cell0 + wid] = running;
        }
    }
    __syncthreads();
    // ---- emission: every survivor to its rank in its cell's list; the score entries go back to zero ----
    for (int k = tid; k < n2; k += FASTG_NT)
    {
        const int y = list2[k] >> 8, x = list2[k] & 255;
        int zs, ze;
        const int c  = cell_of(x, zs, ze);
        const int i  = (y - 3) * (ze - zs) + (x - zs);
        const uint32_t word = bm[c][s_ini[c] ? 1 : 0][i >> 5], bit = 1u << (i & 31);
        const int sc = score[y][x + a];
        score[y][x + a] = 0;
        if (word & bit)
        {
            const int pos = pre[c][i >> 5] + __popc(word & (bit - 1));
            if (pos < P.cell_cap)
                cand[((size_t)img * P.cells_per_img + g.cell0 + c) * P.cell_cap + pos] =
                    (uint32_t)(x + g.off_x) | ((uint32_t)(y + g.off_y) << 12) | ((uint32_t)(sc - 1) << 24);
        }
    }
    g = gnext;
    __syncthreads();  // every read of this group's tile, lists and bitmaps is done before they are reused
    }  // groups of this CTA
}
}  // namespace snk
