// decode_offset_warp.cuh -- the current arg-max stage of the offset-vote decoder for 64-row maps (64x48, 64x64):
// warp-per-band blur, teams of warps, one persistent CTA per SM (decode_offset_team_kernel below).
//
// What limited the pipe kernel (decode_offset_pipe_kernel; ncu, profiles/r02_decode_offset_ncu.txt): 168 registers
// -> 12 warps per SM, each at ~6 cycles per instruction, two block barriers per map (0.94 barrier stalls per issue),
// issue-active 47 %.  Time followed the instruction count, so this version (a) needs no block barrier per map,
// (b) executes fewer instructions, and (c) fits 15 warps per SM:
//
//  * warp = one 16-column band of the map, for BOTH passes.  The row pass always gave a warp the 32 row pairs
//    (p, p + 32) of one 16-column segment; the column pass now walks the same 16 columns (lane = column pair
//    c = lane % 8, 16-row segment g = lane / 8), so it reads only what its own warp wrote: `__syncwarp()`
//    instead of a block barrier between the passes.
//  * T (the row-pass result) has its own buffer, so it never overwrites raw data another warp still reads.
//    T carries the 7 reflected halo rows above and below (written by the lanes that own rows 1..7 and
//    H-8..H-2) and 16 floats of skew after every 16 rows: the 30 loads of a column segment are
//    `[base + constant]` -- no reflection or address arithmetic -- and a warp's 4 segments x 8 column pairs
//    hit each bank twice (256 bytes: the minimum), as do the row pass's stores.
//  * a raw slot (out_i, flipped_i) is handed back to the TMA unit by whichever warp of the team is the last to have
//    copied its window into registers: a shared counter, no waiting (see the kernel's header for the ring).
//  * non-finite maps need no separate path: a warp whose row-pass results are not all finite runs its column
//    pass with np.argmax's NaN rule (arg_better); bands whose results are finite use the plain compare.
// Arithmetic and summation order are those of the pipe kernel: identical bits.
#pragma once
#include "decode_pipe.cuh"

namespace udp {

template <int W, int H, int NT>
struct WarpGeom {
    static constexpr int R = 7;
    using G = PipeGeom<W, H, R, NT>;
    static_assert(H == 64 && G::HP == 32 && G::SEG == 16, "one warp = 32 row pairs x one 16-column band");
    static constexpr int ST = W + 2;         // T row stride (floats; even: column pairs stay 8-byte aligned)
    static constexpr int SKEW = 16;          // floats inserted after every 16 storage rows
    static constexpr int TROWS = H + 2 * R;  // storage row = map row + R
    static_assert((16 * ST) % 32 == 0, "segments 16 rows apart must differ by the skew only");
    __host__ __device__ static constexpr int taddr(int rr) { return rr * ST + (rr >> 4) * SKEW; }
    static constexpr int TFLOATS = ((taddr(TROWS - 1) + W + 31) / 32) * 32;
    static constexpr size_t SMEM = (size_t)(2 * G::BUF + TFLOATS) * sizeof(float);
};

// Row filter from the register window into T (this warp's 16 columns of its 64 rows + the halo copies).
template <int W, int H, int NT>
__device__ __forceinline__ f32x2 warp_row_filter(const f32x2 (&v)[PipeGeom<W, H, 7, NT>::NCOL], float *T, int p, int xs)
{
    constexpr int R = 7;
    using G = PipeGeom<W, H, R, NT>;
    using WG = WarpGeom<W, H, NT>;
    constexpr int PAD = G::PAD, SEG = G::SEG, HP = G::HP;
    const int x0 = xs * SEG;
    float *t0 = T + WG::taddr(p + R) + x0, *t1 = T + WG::taddr(p + HP + R) + x0;
    // reflect-101 halo rows: row -k is row k (k = p), row H-1+k is row H-1-k (k = HP-1-p)
    const bool top = p >= 1 && p <= R, bot = p >= HP - 1 - R && p <= HP - 2;
    float *h0 = T + WG::taddr(top ? R - p : 0) + x0;
    float *h1 = T + WG::taddr(bot ? H - 1 + (HP - 1 - p) + R : 0) + x0;
    f32x2 acc = pk(0.f, 0.f);  // sum of everything written: finite iff every result is (up to overflow of the sum)
#pragma unroll
    for (int o = 0; o < SEG; o += 2) {
        f32x2 s0 = fmul2p(v[PAD - R + o], Taps<R>::k(0)), s1 = fmul2p(v[PAD - R + o + 1], Taps<R>::k(0));
#pragma unroll
        for (int i = 1; i < 2 * R + 1; ++i) {
            s0 = ffma2p(v[PAD - R + o + i], Taps<R>::k(i), s0);
            s1 = ffma2p(v[PAD - R + o + 1 + i], Taps<R>::k(i), s1);
        }
        const f32x2 ra = pk(lo_of(s0), lo_of(s1)), rb = pk(hi_of(s0), hi_of(s1));
        *reinterpret_cast<f32x2 *>(t0 + o) = ra;
        *reinterpret_cast<f32x2 *>(t1 + o) = rb;
        if (top) *reinterpret_cast<f32x2 *>(h0 + o) = ra;
        if (bot) *reinterpret_cast<f32x2 *>(h1 + o) = rb;
        acc = fadd2p(acc, fadd2p(s0, s1));
    }
    return acc;
}

// Column pass over this warp's band: lane = (column pair c, 16-row segment g); emit(y, x, a, b) gets the blurred
// values a = (y, x), b = (y, x + 1), rows ascending.
template <int W, int H, int NT, typename Emit>
__device__ __forceinline__ void warp_col_pass(const float *T, int warp, int lane, Emit emit)
{
    constexpr int R = 7, RS = 16;
    using WG = WarpGeom<W, H, NT>;
    const int c = lane & 7, g = lane >> 3;
    const float *base = T + WG::taddr(RS * g) + 16 * warp + 2 * c;  // storage row 16 g = map row 16 g - R
    f32x2 t[RS + 2 * R];
#pragma unroll
    for (int i = 0; i < RS + 2 * R; ++i) t[i] = *reinterpret_cast<const f32x2 *>(base + i * WG::ST + (i >> 4) * WG::SKEW);
#pragma unroll
    for (int o = 0; o < RS; ++o) {
        f32x2 s = ffma2p(t[o + R], Taps<R>::k(R), pk(0.f, 0.f));
#pragma unroll
        for (int i = 1; i <= R; ++i) s = ffma2p(fadd2p(t[o + R + i], t[o + R - i]), Taps<R>::k(R + i), s);
        emit(RS * g + o, 16 * warp + 2 * c, lo_of(s), hi_of(s));
    }
}

// ---------------------------------------------------------------------------------------
// ONE persistent CTA per SM made of NTEAM independent teams (a team = the NW warps of one map) that share a RING of
// NP raw (out, flipped) slots.
//
// Why not one team per CTA with its own raw pair (measured: 1.72 ms, profiles/r02_dec_warp_v1): a team's next map is
// requested only when its slowest warp has copied its window, and needed one blur later -- less than a loaded HBM
// round trip, so 15 % of the warp samples sat in the mbarrier wait; a second pair per CTA does not fit four CTAs.
// Pooled in one CTA, the slots go to whichever map is next: the CTA's maps k = 0, 1, 2, ... land in slot k % NP in
// order, team t takes k = t, t + NTEAM, ...; a slot is refilled (with map k + NP) by the last warp of its team to
// have copied its window.  (With NP == NTEAM that would tie each team to one slot again; the deal is then rotated.)
// 128 registers without spills give five teams = 15 warps per SM at 64x48 (1.72 -> 1.56 ms per 16 384 persons x 17
// maps; four teams of 168 registers: 1.78 ms, six of 96: 1.77 ms).
// No block barrier after the prologue: a team meets on its named barrier (bar.sync team+1) only to vote.
// ---------------------------------------------------------------------------------------
template <int W, int H, int TNT, int NTEAM, int NP, bool FLIP>
__global__ void __launch_bounds__(TNT * NTEAM, 1)
decode_offset_team_kernel(const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ CUtensorMap tm_fl,
                          DecodeOffsetArgs p, long long maps)
{
    constexpr int R = 7;
    using G = PipeGeom<W, H, R, TNT>;
    using WG = WarpGeom<W, H, TNT>;
    // Shared memory is budgeted to stay below 196 KiB per SM (ring + T buffers + mailboxes + the 1 KiB the system keeps):
    // one step further up the carve-out ladder leaves 28 KiB of L1 instead of 60 and costs 15 % (1.52 -> 1.75 ms,
    // measured by padding the allocation) -- the vote's scattered loads live on L1 hits.  That is why a fifth slot
    // measured slower; the mailboxes take what is left (24 maps at 64x48 with the mirrored partner: 2 % over 16).
    constexpr int NW = G::NWARP;
    constexpr int B_V = 0, B_I = NW, B_MAP = 2 * NW, BOXW = 2 * NW + 1;
    constexpr int SLOT = (FLIP ? 2 : 1) * G::BUF;  // floats per ring slot: out, then flipped
    constexpr int DYN_BYTES = (NP * SLOT + NTEAM * WG::TFLOATS) * 4 + 128;
    constexpr int BOX_BYTES = NTEAM * (2 * BOXW + 14) * 4;  // one more map in every team's two mailboxes and row buffer
    constexpr int ROOM = (196 * 1024 - 1024 - DYN_BYTES - 256) / BOX_BYTES;
    constexpr int NBOX = ROOM >= 32 ? 32 : (ROOM & ~3);  // (one map per lane in the vote's last step: at most 32)
    static_assert(NBOX >= 8, "shared memory budget");
    static_assert(TNT == 32 * NW && NW == G::XS, "whole warps, one per column band");
    // Which team takes the CTA's k-th map, k = NP * f + s (slot s, fill f).  NP != NTEAM: team k % NTEAM (the two must be
    // coprime, or a team would only ever see some of the slots).  NP == NTEAM: team (s + f) % NTEAM -- every group of NP
    // maps is dealt round the teams one place further on, so team t's f-th map sits in slot (t - f) mod NP.
    constexpr bool ROT = NP == NTEAM;
    static_assert(ROT || (NP % NTEAM != 0 && NTEAM % NP != 0), "slots and teams");
    constexpr int ITEMS = (NBOX * 14 + TNT - 1) / TNT;  // vote items per thread
    constexpr unsigned SLOT_BYTES = (FLIP ? 2u : 1u) * (unsigned)(G::SR * H * 4);
    // dynamic shared memory is only guaranteed 16-byte alignment; TMA destinations need 128
    extern __shared__ __align__(16) float smem_raw[];
    float *smem = smem_raw + ((128u - ((unsigned)__cvta_generic_to_shared(smem_raw) & 127u)) & 127u) / 4u;  // still a shared-space pointer
    __shared__ __align__(8) unsigned long long s_bars[NP];
    __shared__ unsigned s_done[NP];                       // warps that have copied their window of the slot's map
    __shared__ int s_fill[NP];                            // which fill the slot's barrier has been armed for
    __shared__ float s_box[NTEAM][2 * NBOX * BOXW];       // per team, two mailboxes: one fills while the other is voted on
    __shared__ float s_rows[NTEAM][NBOX * 14];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int team = warp / NW, tw = warp - team * NW, ttid = tid - team * TNT;
    float *T = smem + NP * SLOT + team * WG::TFLOATS;
    const unsigned ring0 = (unsigned)__cvta_generic_to_shared(smem), bars = (unsigned)__cvta_generic_to_shared(s_bars);
    // this CTA's maps: m(k) = blockIdx.x + k * gridDim.x, k < KC (maps < 2^31, checked by the launcher)
    const int KC = (int)((maps - (long long)blockIdx.x + (long long)gridDim.x - 1) / (long long)gridDim.x);

    auto issue = [&](int s, int f) {  // one thread: the loads of fill f of slot s, i.e. of this CTA's map k = NP * f + s
        const unsigned m = blockIdx.x + (unsigned)(NP * f + s) * gridDim.x;
        const unsigned bar = bars + 8u * (unsigned)s, dst = ring0 + (unsigned)s * (unsigned)(SLOT * 4);
        *reinterpret_cast<volatile int *>(&s_fill[s]) = f;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(SLOT_BYTES) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(dst), "l"((unsigned long long)(uintptr_t)&tm_out), "r"(0), "r"((int)(m * 3u * H)), "r"(bar) : "memory");
        if (FLIP) {
            const unsigned n = m / (unsigned)p.K, kj = m - n * (unsigned)p.K;
            const int row = (int)((n * (unsigned)p.K + (unsigned)__ldg(p.flip_perm + kj)) * 3u * H);
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(dst + (unsigned)(G::BUF * 4)), "l"((unsigned long long)(uintptr_t)&tm_fl), "r"(0), "r"(row), "r"(bar) : "memory");
        }
    };
    auto team_bar = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "n"(TNT) : "memory"); };
    auto entry = [&](const float *b, float &bv, int &bi) -> long long {  // combine the per-warp candidates
        bv = b[B_V];
        bi = __float_as_int(b[B_I]);
#pragma unroll
        for (int q = 1; q < NW; ++q) {
            const float v = b[B_V + q];
            const int i = __float_as_int(b[B_I + q]);
            if (arg_better(v, i, bv, bi)) { bv = v; bi = i; }
        }
        return (long long)__float_as_int(b[B_MAP]);
    };
    auto flush = [&](int count, int half) {  // the team votes on the `count` maps of its mailbox `half`
        const float *boxes = s_box[team] + half * NBOX * BOXW;
        float *rows = s_rows[team];
        team_bar();  // every warp of the team has posted its candidates (and is done with the previous vote's rows)
        // (map, channel, row) items: unrolled, so that the scattered global loads of all of a thread's items are in
        // flight together (rolled, the ITEMS round trips to HBM come one after the other)
        float res[ITEMS];
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
            const int item = ttid + it * TNT;
            res[it] = 0.0f;
            if (item < count * 14) {
                const int e = item / 14, r = item - e * 14;
                float bv;
                int bi;
                const long long m = entry(boxes + e * BOXW, bv, bi);
                res[it] = offset_row_item(p, m, bv, bi, r / 7, r - (r / 7) * 7);
            }
        }
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
            const int item = ttid + it * TNT;
            if (item < count * 14) rows[item] = res[it];
        }
        team_bar();
        if (tw == 0 && lane < count) {  // one map per lane; the other warps are already filling the other mailbox
            float bv;
            int bi;
            const long long m = entry(boxes + lane * BOXW, bv, bi);
            offset_final(p, m, bv, bi, rows + lane * 14);
        }
    };

    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8 * i) : "memory");
            s_done[i] = 0u;
            s_fill[i] = -1;
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0)
        for (int s = 0; s < NP && s < KC; ++s) issue(s, 0);

    // this team's sequence of (slot, fill): see ROT above
    int s = ROT ? team : team % NP, f = ROT ? 0 : team / NP;
    int slot_n = 0, half = 0;
    while (NP * f + s < KC) {
        const int k = NP * f + s;
        if (slot_n == NBOX) { flush(NBOX, half); slot_n = 0; half ^= 1; }
        // A parity wait can only tell "this phase" from "the one before": a team that is more than a fill ahead of the team
        // that frees its next slot (the slot's barrier still in the phase before last) would be let through.  So first
        // see the slot armed for OUR fill (a plain flag written by the issuer before it arms the barrier).
        while (*reinterpret_cast<volatile int *>(&s_fill[s]) != f) {}
        mbar_wait(bars + 8u * (unsigned)s, (unsigned)(f & 1));  // per-thread wait on the slot's fill
        const float *bufP = smem + s * SLOT, *bufQ = bufP + G::BUF;

        f32x2 v[G::NCOL];
        pipe_window<W, H, R, TNT, FLIP, false>(bufP, bufQ, lane, tw, v);
        // every lane's window is in registers: this warp is done with the slot; the last warp of the team to say so
        // refills it with the map NP places further on
        __syncwarp();
        if (lane == 0) {
            if (atomicAdd(&s_done[s], 1u) == (unsigned)(NW - 1)) {
                s_done[s] = 0u;
                if (k + NP < KC) issue(s, f + 1);
            }
        }
        const f32x2 tsum = warp_row_filter<W, H, TNT>(v, T, lane, tw);
        const bool wbad = __any_sync(0xffffffffu, !(fabsf(lo_of(tsum) + hi_of(tsum)) < INFINITY));
        __syncwarp();  // this warp's band of T is complete
        float bv;
        int bi;
        const int xc = 16 * tw + 2 * (lane & 7);
        if (!wbad) {
            float v0 = -INFINITY, v1 = -INFINITY;
            int y0 = 0, y1 = 0;
            warp_col_pass<W, H, TNT>(T, tw, lane, [&](int y, int, float a, float b) {
                if (a > v0) { v0 = a; y0 = y; }
                if (b > v1) { v1 = b; y1 = y; }
            });
            bv = v0;
            bi = y0 * W + xc;
            if (arg_better(v1, y1 * W + xc + 1, bv, bi)) { bv = v1; bi = y1 * W + xc + 1; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {  // (also orders this map's reads of T before the next map's stores)
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
        } else {  // NaN / inf in this band's row-pass results: np.argmax's rule (the first NaN wins)
            bv = -INFINITY;
            bi = 0x7fffffff;
            warp_col_pass<W, H, TNT>(T, tw, lane, [&](int y, int x, float a, float b) {
                if (arg_better(a, y * W + x, bv, bi)) { bv = a; bi = y * W + x; }
                if (arg_better(b, y * W + x + 1, bv, bi)) { bv = b; bi = y * W + x + 1; }
            });
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (arg_better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
            }
        }
        float *box = s_box[team] + (half * NBOX + slot_n) * BOXW;
        if (lane == 0) { box[B_V + tw] = bv; box[B_I + tw] = __int_as_float(bi); }
        if (ttid == 0) box[B_MAP] = __int_as_float((int)(blockIdx.x + (unsigned)k * gridDim.x));
        ++slot_n;
        if (ROT) {  // next group, one slot back
            s = s == 0 ? NP - 1 : s - 1;
            ++f;
        } else {  // k += NTEAM
            s += NTEAM % NP;
            f += NTEAM / NP;
            if (s >= NP) { s -= NP; ++f; }
        }
    }
    if (slot_n) flush(slot_n, half);
}

}  // namespace udp
