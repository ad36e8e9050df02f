// k_describe with its two patches staged by TMA (included by orb_kernels.cuh; the arithmetic is k_describe's, line by line).
//
// k_describe is bound by the L1 data pipe (87 % of its wavefront peak, profiles/r2n_ncu_summary.csv): per keypoint 279
// word loads of the level image for the moments and 512 byte gathers of the blurred level for rBRIEF, all through global
// memory. Here a warp fetches, per keypoint, the 31 rows x 48 bytes around the moment disc and the 39 rows x 64 bytes that
// hold every rotated sample (pattern radius 18.4, keypoints keep >= 19 pixels from the image's sides) as two tensor boxes
// (start rounded down to 16 bytes) into its own shared-memory slots, two keypoints ahead, completion on per-slot mbarriers;
// the loads of the arithmetic become shared-memory loads, which cost the data pipe one wavefront per conflict-free
// instruction instead of one per touched line.
#pragma once

namespace snk
{
constexpr int DT_IMG_W = 48, DT_IMG_H = 31, DT_IMG_BYTES = 1536;  // 48 x 31 = 1488, slots 128-byte aligned
constexpr int DT_G_W = 64, DT_G_H = 39, DT_G_BYTES = 2560;        // 64 x 39 = 2496
// per warp two slots, each holding first a moment patch and later a blurred patch: the two phases of a warp's batch do not
// overlap, and shared memory is what limits the resident warps (8 KB per warp: 24 warps per SM, 0.548 ms; with 16 warps
// 0.670 ms) — a slot takes its first blurred patch as soon as its last moment patch has been read
constexpr int DT_WARP_BYTES = 2 * DT_G_BYTES;
constexpr int DT_SMEM = DS_WARPS * DT_WARP_BYTES;
struct DescMaps
{
    CUtensorMap img[8];    // per level: the level image, box DT_IMG_W x DT_IMG_H x 1
    CUtensorMap gauss[8];  // per level: the blurred level, box DT_G_W x DT_G_H x 1
};

__global__ void __launch_bounds__(DS_WARPS * 32) k_describe_tma(PyrDev P, const __grid_constant__ DescMaps maps,
                                                                const uint32_t* __restrict__ sel, const int* __restrict__ sel_count,
                                                                snk_keypoint_f32* __restrict__ kps, uint64_t* __restrict__ desc,
                                                                int* __restrict__ counts, const float2* __restrict__ pattern_t,
                                                                const uint2* __restrict__ moment_tab, int* __restrict__ err)
{
    extern __shared__ __align__(128) unsigned char dt_smem[];
    __shared__ float2 pat[16][32];  // pat[j][lane] = point 16*lane + j as floats (conflict-free: lanes are consecutive)
    __shared__ int s_pref[SNK_MAX_LEVELS + 1];
    __shared__ __align__(8) unsigned long long s_bar[DS_WARPS][4];  // per warp: image slots 0 / 1, blurred slots 0 / 1
    const int img = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int i = tid; i < 512; i += DS_WARPS * 32) (&pat[0][0])[i] = __ldg(pattern_t + i);
    if (tid == 0)
    {
        int acc = 0;
        for (int l = 0; l < P.nlevels; ++l)
        {
            s_pref[l] = acc;
            acc += sel_count[img * P.nlevels + l];
        }
        s_pref[P.nlevels] = acc;
        if (blockIdx.x == 0) counts[img] = min(acc, P.kp_stride);
    }
    if (lane == 0)
    {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&s_bar[wid][i])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int total = min(s_pref[P.nlevels], P.kp_stride);
    // the warp's keypoints: g(kk) = (blockIdx.x * DS_KP_PER_WARP + kk) * DS_WARPS + wid, kk < nk
    const int g0 = blockIdx.x * DS_KP_PER_WARP * DS_WARPS + wid;
    if (g0 >= total) return;
    const int nk = min(DS_KP_PER_WARP, (total - g0 + DS_WARPS - 1) / DS_WARPS);

    // ---- per keypoint (lane kk owns keypoint kk): level and packed position ----
    int my_level = 0;
    uint32_t my_p = 0;
    if (lane < nk)
    {
        const int g = g0 + lane * DS_WARPS;
        while (my_level + 1 < P.nlevels && g >= s_pref[my_level + 1]) ++my_level;
        my_p = sel[(size_t)img * P.sel_per_img + P.lv[my_level].sel_base + (g - s_pref[my_level])];
    }
    unsigned char* slots   = dt_smem + wid * DT_WARP_BYTES;
    const unsigned slots_a = (unsigned)__cvta_generic_to_shared(slots);
    const unsigned bar_a   = (unsigned)__cvta_generic_to_shared(&s_bar[wid][0]);
    // the boxes of keypoint kk into slot kk & 1 (all lanes call: the shuffles are warp-wide; lane 0 issues)
    auto fetch = [&](int kk, bool blurred) {
        const int level  = __shfl_sync(0xffffffffu, my_level, kk);
        const uint32_t p = __shfl_sync(0xffffffffu, my_p, kk);
        if (lane == 0)
        {
            const int x = (int)(p & 0xfffu) + 16, y = (int)((p >> 12) & 0xfffu) + 16;
            const unsigned bar = bar_a + 8u * ((blurred ? 2 : 0) + (kk & 1));
            const unsigned dst = slots_a + (unsigned)((kk & 1) * DT_G_BYTES);
            const CUtensorMap* m = blurred ? &maps.gauss[level] : &maps.img[level];
            const int bx = blurred ? ((x - 19) & ~15) : ((x - 15) & ~15), by = blurred ? y - 19 : y - 15;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
                         "r"((unsigned)(blurred ? DT_G_W * DT_G_H : DT_IMG_W * DT_IMG_H))
                         : "memory");
            asm volatile(
                "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
                "l"(reinterpret_cast<unsigned long long>(m)), "r"(bx), "r"(by), "r"(img), "r"(bar)
                : "memory");
        }
    };
    bool ok = true;
    auto wait = [&](int kk, bool blurred) {
        const unsigned bar = bar_a + 8u * ((blurred ? 2 : 0) + (kk & 1)), parity = (unsigned)((kk >> 1) & 1);
        unsigned done = 0;
        for (int spin = 0; !done && spin < (1 << 22); ++spin)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x989680;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done)
                         : "r"(bar), "r"(parity)
                         : "memory");
        ok = ok && done;
    };
    fetch(0, false);
    if (nk > 1) fetch(1, false);

    // ---- ORB::ComputeAngle: integer moments over the radius-15 disc of the un-blurred level ----
    int my_m10 = 0, my_m01 = 0;
#pragma unroll 1
    for (int kk = 0; kk < nk; ++kk)
    {
        const uint32_t p = __shfl_sync(0xffffffffu, my_p, kk);
        const int x      = (int)(p & 0xfffu) + 16;
        const int al     = (x - 15) & 3;
        // word 0 of the 36-byte window inside the box: the box starts at the 16-byte boundary below x - 15
        const unsigned* base = reinterpret_cast<const unsigned*>(slots + (kk & 1) * DT_G_BYTES) + (((x - 15 - al) - ((x - 15) & ~15)) >> 2);
        const uint2* tab     = moment_tab + al * DS_MOM_TASKS;
        wait(kk, false);
        int m10 = 0, m01 = 0;
#pragma unroll
        for (int it = 0; it < (DS_MOM_TASKS + 31) / 32; ++it)
        {
            const int t = lane + 32 * it;
            if (t < DS_MOM_TASKS)
            {
                const int r = (t * 7282) >> 16;  // t / 9 for t < 279
                const int w = t - 9 * r;
                const unsigned word = base[r * (DT_IMG_W / 4) + w];
                const uint2 e       = __ldg(tab + t);
                m10 = dp4a_u8s8(word, e.x, m10);               // u8 pixels x s8 column offsets
                m01 += (r - 15) * (int)__dp4a(word, e.y, 0u);  // v * sum of the row's pixels inside the disc
            }
        }
        m10 = warp_sum(m10);
        m01 = warp_sum(m01);
        if (lane == kk) my_m10 = m10, my_m01 = m01;
        __syncwarp();  // the slot is read: the keypoint after next may land in it
        if (kk + 2 < nk)
            fetch(kk + 2, false);
        else
            fetch(kk & 1, true);  // the slot's last moment patch: the blurred patch of keypoint 0 or 1 arrives during the trigonometry
    }

    // ---- angle, sine and cosine of all keypoints of the warp at once (lane kk computes keypoint kk) ----
    float my_angle = snkmath::atan2f_fdlibm((float)my_m01, (float)my_m10);
    my_angle       = __fdiv_rn(__fmul_rn(my_angle, 180.0f), 3.14159265358979323846f);
    my_angle       = __fadd_rn((my_angle < 0.f) ? 360.f : 0.f, my_angle);
    const float my_rad = __fmul_rn(__fdiv_rn(my_angle, 180.0f), 3.14159265358979323846f);
    const float my_a = snkmath::cosf_glibc(my_rad), my_b = snkmath::sinf_glibc(my_rad);

    // ---- ORB::ComputeDescriptor on the blurred level; lane i produces byte i ----
#pragma unroll 1
    for (int kk = 0; kk < nk; ++kk)
    {
        const int g       = g0 + kk * DS_WARPS;
        const int level   = __shfl_sync(0xffffffffu, my_level, kk);
        const uint32_t p  = __shfl_sync(0xffffffffu, my_p, kk);
        const float angle = __shfl_sync(0xffffffffu, my_angle, kk);
        const float a     = __shfl_sync(0xffffffffu, my_a, kk);
        const float b     = __shfl_sync(0xffffffffu, my_b, kk);
        const LevelDev& L = P.lv[level];
        const int x = (int)(p & 0xfffu) + 16, y = (int)((p >> 12) & 0xfffu) + 16;
        const float resp = (float)(p >> 24);
        // sample (ix, iy) of the level lies at [iy - (y - 19)][ix - box start] of the slot
        const uint8_t* gs = slots + (kk & 1) * DT_G_BYTES - ((y - 19) * DT_G_W + ((x - 19) & ~15));
        const float px = (float)x, py = (float)y;
        wait(kk, true);
        uint32_t byte = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j)
        {
            int tv[2];
#pragma unroll
            for (int e = 0; e < 2; ++e)
            {
                const float2 q = pat[2 * j + e][lane];
                const float fx = __fadd_rn(px, __fsub_rn(__fmul_rn(q.x, a), __fmul_rn(q.y, b)));
                const float fy = __fadd_rn(py, __fadd_rn(__fmul_rn(q.x, b), __fmul_rn(q.y, a)));
                // Saiga::iRound: (int)(v + (v >= 0 ? 0.5f : -0.5f)); the half takes the sign bit of v
                const int ix = (int)__fadd_rn(fx, __int_as_float(0x3f000000 | (__float_as_int(fx) & 0x80000000)));
                const int iy = (int)__fadd_rn(fy, __int_as_float(0x3f000000 | (__float_as_int(fy) & 0x80000000)));
                tv[e]        = gs[iy * DT_G_W + ix];
            }
            byte |= (uint32_t)(tv[0] < tv[1]) << j;
        }
        __syncwarp();
        if (kk + 2 < nk) fetch(kk + 2, true);
        // 4 lanes -> one little-endian 32-bit word
        uint32_t word = byte << (8 * (lane & 3));
        word |= __shfl_xor_sync(0xffffffffu, word, 1);
        word |= __shfl_xor_sync(0xffffffffu, word, 2);
        const size_t o = (size_t)img * P.kp_stride + g;
        if ((lane & 3) == 0) reinterpret_cast<uint32_t*>(desc + 4 * o)[lane >> 2] = word;

        // ---- output record (ORBExtractor.cpp:163-171, 223-232) ----
        if (lane < 6)
        {
            float f;
            const float sc = L.scale;
            if (lane == 0)
                f = (level != 0) ? __fmul_rn(px, sc) : px;
            else if (lane == 1)
                f = (level != 0) ? __fmul_rn(py, sc) : py;
            else if (lane == 2)
                f = L.size;
            else if (lane == 3)
                f = angle;
            else if (lane == 4)
                f = resp;
            else
                f = __int_as_float(level);
            reinterpret_cast<float*>(kps + o)[lane] = f;
        }
    }  // keypoints of this warp
    if (!ok && lane == 0) atomicExch(err, 4);
}
}  // namespace snk
