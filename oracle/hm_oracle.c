/*
 * hm_oracle.c — TEST INFRASTRUCTURE (see hm_oracle.h).  Plain-C restatement of the
 * reference's plane encoder, policy index, masked softmax and value shaping.
 * Every function cites the reference file:line it follows.
 */
#include "hm_oracle.h"

#include <math.h>
#include <string.h>

/* ------------------------------------------------------------------------------------
 * IEEE helpers
 * ---------------------------------------------------------------------------------- */
static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

uint16_t hmo_float_to_half(float f) {
    uint32_t x = f2u(f);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t ax = x & 0x7fffffffu;
    if (ax >= 0x7f800000u) { /* inf / nan */
        return (uint16_t)(sign | 0x7c00u | ((ax > 0x7f800000u) ? 0x0200u : 0u));
    }
    if (ax >= 0x477ff000u) { /* rounds to >= 65520 -> inf */
        return (uint16_t)(sign | 0x7c00u);
    }
    if (ax < 0x38800000u) { /* subnormal half or zero: value < 2^-14 */
        if (ax < 0x33000000u) return (uint16_t)sign; /* < 2^-25 -> 0 (ties-to-even at 2^-25 -> 0) */
        uint32_t exp = ax >> 23;
        uint32_t man = (ax & 0x7fffffu) | 0x800000u;
        uint32_t shift = 126u - exp; /* 14..24: man is 1.23 fixed point; want units of 2^-24 */
        uint32_t half_man = man >> shift;
        uint32_t rem = man & ((1u << shift) - 1u);
        uint32_t halfway = 1u << (shift - 1u);
        if (rem > halfway || (rem == halfway && (half_man & 1u))) half_man++;
        return (uint16_t)(sign | half_man);
    }
    {
        uint32_t exp = (ax >> 23) - 112u;
        uint32_t man = ax & 0x7fffffu;
        uint32_t h = (exp << 10) | (man >> 13);
        uint32_t rem = man & 0x1fffu;
        if (rem > 0x1000u || (rem == 0x1000u && (h & 1u))) h++;
        return (uint16_t)(sign | h);
    }
}

float hmo_half_to_float(uint16_t h) {
    uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu;
    uint32_t man = h & 0x3ffu;
    if (exp == 0) {
        if (man == 0) return u2f(sign);
        /* subnormal: man * 2^-24 */
        float v = (float)man * 5.9604644775390625e-08f;
        return (sign ? -v : v);
    }
    if (exp == 31) return u2f(sign | 0x7f800000u | (man << 13));
    return u2f(sign | ((exp + 112u) << 23) | (man << 13));
}

uint16_t hmo_float_to_bf16(float f) {
    uint32_t x = f2u(f);
    if ((x & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((x >> 16) | 0x0040u);
    uint32_t lsb = (x >> 16) & 1u;
    x += 0x7fffu + lsb;
    return (uint16_t)(x >> 16);
}
float hmo_bf16_to_float(uint16_t h) { return u2f((uint32_t)h << 16); }

/* ------------------------------------------------------------------------------------
 * Stage P — planes (planes.cc)
 * ---------------------------------------------------------------------------------- */
static uint64_t bswap64(uint64_t x) { return __builtin_bswap64(x); } /* utils.h:111-113 */

/* planes.cc:31-38 set_bits_from_bitmap */
static void plane_bits(float* plane, uint64_t bb) {
    while (bb) {
        int idx = __builtin_ctzll(bb);
        plane[idx] = 1.0f;
        bb &= bb - 1;
    }
}
/* planes.cc:42-63 fill_plane */
static void plane_fill(float* plane, float v) {
    for (int i = 0; i < 64; ++i) plane[i] = v;
}

/* One board's 37 channels: planes.cc:236-245 (board 0) / 247-256 (board 1). */
static void encode_board(const hm_packed_board* b, int board_idx, int team_black, int tadv,
                         float* out /* [37][64], pre-zeroed */) {
    /* planes.cc:90-93 needs_flipping: board A iff team BLACK, board B iff team WHITE */
    const int flip = (board_idx == 0) ? team_black : !team_black;
    /* planes.cc:99-100: first = team on A, ~team on B */
    const int first = (board_idx == 0) ? team_black : !team_black;
    const int second = !first;
    const int order[2] = {first, second};
    float* p = out;

    /* ch 0-11 pieces — planes.cc:97-111 */
    for (int ci = 0; ci < 2; ++ci)
        for (int pt = 0; pt < 6; ++pt) {
            uint64_t bb = b->pieces[order[ci] * 6 + pt];
            if (flip) bb = bswap64(bb);
            plane_bits(p, bb);
            p += 64;
        }
    /* ch 12-21 pockets — planes.cc:114-124: (float)count / 16.0f */
    for (int ci = 0; ci < 2; ++ci)
        for (int pt = 0; pt < 5; ++pt) {
            plane_fill(p, (float)b->pocket[order[ci] * 5 + pt] / 16.0f);
            p += 64;
        }
    /* ch 22-23 promoted — planes.cc:127-138 */
    for (int ci = 0; ci < 2; ++ci) {
        uint64_t bb = b->promoted[order[ci]];
        if (flip) bb = bswap64(bb);
        plane_bits(p, bb);
        p += 64;
    }
    /* ch 24 en passant — planes.cc:141-151 */
    {
        uint64_t bb = (b->ep_square < 64) ? (1ULL << b->ep_square) : 0ULL;
        if (flip) bb = bswap64(bb);
        plane_bits(p, bb);
        p += 64;
    }
    /* ch 25 colour info — planes.cc:154-159: stm == expected (team on A, ~team on B) */
    {
        const int stm_black = (b->flags & HM_FLAG_STM_BLACK) != 0;
        plane_fill(p, (stm_black == first) ? 1.0f : 0.0f);
        p += 64;
    }
    /* ch 26 constant ones — planes.cc:242,253 */
    plane_fill(p, 1.0f);
    p += 64;
    /* ch 27-30 castling — planes.cc:162-176.  On both boards the order is
     * {first OO, first OOO, second OO, second OOO} with first as above. */
    {
        const int oo[2] = {HM_FLAG_WHITE_OO, HM_FLAG_BLACK_OO};
        const int ooo[2] = {HM_FLAG_WHITE_OOO, HM_FLAG_BLACK_OOO};
        plane_fill(p, (b->flags & oo[first]) ? 1.0f : 0.0f);  p += 64;
        plane_fill(p, (b->flags & ooo[first]) ? 1.0f : 0.0f); p += 64;
        plane_fill(p, (b->flags & oo[second]) ? 1.0f : 0.0f); p += 64;
        plane_fill(p, (b->flags & ooo[second]) ? 1.0f : 0.0f); p += 64;
    }
    /* ch 31 time advantage — planes.cc:244,255 */
    plane_fill(p, tadv ? 1.0f : 0.0f);
    p += 64;
    /* ch 32-33 last move — planes.cc:183-198; orient_square = sq ^ 56 (:178-180) */
    if (b->last_to < 64) {
        if (b->last_from < 64) p[flip ? (b->last_from ^ 56) : b->last_from] = 1.0f;
        p += 64;
        p[flip ? (b->last_to ^ 56) : b->last_to] = 1.0f;
        p += 64;
    } else {
        p += 128;
    }
    /* ch 34 halfmove clock — planes.cc:203-206: min(rule50, 50) / 50.0f */
    {
        int r = b->rule50 < 50 ? b->rule50 : 50;
        plane_fill(p, (float)r / 50.0f);
        p += 64;
    }
    /* ch 35-36 repetitions — planes.cc:208-210 */
    plane_fill(p, b->repetitions >= 2 ? 1.0f : 0.0f); p += 64;
    plane_fill(p, b->repetitions >= 3 ? 1.0f : 0.0f);
}

void hmo_encode_planes_f32(const hm_packed_pos* pos, size_t n, float* out) {
    for (size_t i = 0; i < n; ++i) {
        float* o = out + i * HM_NB_INPUT_VALUES;
        memset(o, 0, sizeof(float) * HM_NB_INPUT_VALUES); /* planes.cc:215-233 */
        const int team_black = (pos[i].board[0].flags & HM_FLAG_TEAM_BLACK) != 0;
        const int tadv = (pos[i].board[0].flags & HM_FLAG_TIME_ADVANTAGE) != 0;
        encode_board(&pos[i].board[0], 0, team_black, tadv, o);
        encode_board(&pos[i].board[1], 1, team_black, tadv, o + 37 * 64);
    }
}

void hmo_encode_planes_f16(const hm_packed_pos* pos, size_t n, uint16_t* out) {
    float tmp[HM_NB_INPUT_VALUES];
    for (size_t i = 0; i < n; ++i) {
        hmo_encode_planes_f32(pos + i, 1, tmp);
        for (int k = 0; k < HM_NB_INPUT_VALUES; ++k)
            out[i * HM_NB_INPUT_VALUES + k] = hmo_float_to_half(tmp[k]); /* planes.cc:22-28 */
    }
}

void hmo_encode_planes_bf16(const hm_packed_pos* pos, size_t n, uint16_t* out) {
    float tmp[HM_NB_INPUT_VALUES];
    for (size_t i = 0; i < n; ++i) {
        hmo_encode_planes_f32(pos + i, 1, tmp);
        for (int k = 0; k < HM_NB_INPUT_VALUES; ++k)
            out[i * HM_NB_INPUT_VALUES + k] = hmo_float_to_bf16(tmp[k]);
    }
}

void hmo_encode_planes_u8(const hm_packed_pos* pos, size_t n, uint8_t* out) {
    float tmp[HM_NB_INPUT_VALUES];
    for (size_t i = 0; i < n; ++i) {
        hmo_encode_planes_f32(pos + i, 1, tmp);
        for (int k = 0; k < HM_NB_INPUT_VALUES; ++k) {
            float v = tmp[k] < 0.0f ? 0.0f : (tmp[k] > 1.0f ? 1.0f : tmp[k]);
            out[i * HM_NB_INPUT_VALUES + k] = (uint8_t)lroundf(v * 255.0f); /* selfplay.cc:471-473 */
        }
    }
}

/* ------------------------------------------------------------------------------------
 * Stage D — policy index (globals.cc:82-136, constants.h:24)
 *
 * Label layout of UCI_MOVES[4672], index = channel*64 + square (from-square for board
 * moves, to-square for drops), squares a1 = 0:
 *   ch 0        "pass" in all 64 slots; POLICY_INDEX keeps the FIRST occurrence -> index 0
 *   ch 1        P@sq for ranks 2..7 only
 *   ch 2..5     N@, R@, B@, Q@ (note R before B)
 *   ch 6..61    queen-like rays: direction d in {N,NE,E,SE,S,SW,W,NW}, distance k in 1..7,
 *               channel 6 + 7*d + (k-1), label from+to when `to` is on the board
 *   ch 62..69   knight jumps (df,dr) = (1,2),(2,1),(2,-1),(1,-2),(-1,-2),(-2,-1),(-2,1),(-1,2)
 *   ch 70..72   knight under-promotion from rank 7: capture-left, push, capture-right ("..n")
 * Colour 1 (black to move) looks the move up after mirroring ranks (sq ^ 56),
 * globals.cc:36-52 fast_mirror_uci.
 * ---------------------------------------------------------------------------------- */
static int32_t g_tab_normal[2][64][64][2];
static int32_t g_tab_drop[2][64][8];
static int g_policy_ready = 0;

static int label_index_normal(int from, int to) { /* index of label "from+to" or -1 */
    static const int ddf[8] = {0, 1, 1, 1, 0, -1, -1, -1};
    static const int ddr[8] = {1, 1, 0, -1, -1, -1, 0, 1};
    static const int kdf[8] = {1, 2, 2, 1, -1, -2, -2, -1};
    static const int kdr[8] = {2, 1, -1, -2, -2, -1, 1, 2};
    const int ff = from & 7, fr = from >> 3, tf = to & 7, tr = to >> 3;
    const int df = tf - ff, dr = tr - fr;
    if (df == 0 && dr == 0) return -1;
    for (int d = 0; d < 8; ++d)
        for (int k = 1; k <= 7; ++k)
            if (df == ddf[d] * k && dr == ddr[d] * k) return (6 + 7 * d + (k - 1)) * 64 + from;
    for (int j = 0; j < 8; ++j)
        if (df == kdf[j] && dr == kdr[j]) return (62 + j) * 64 + from;
    return -1;
}
static int label_index_knight_promo(int from, int to) { /* label "from+to+n" */
    const int ff = from & 7, fr = from >> 3, tf = to & 7, tr = to >> 3;
    if (fr != 6 || tr != 7) return -1;
    const int df = tf - ff;
    if (df < -1 || df > 1) return -1;
    return (71 + df) * 64 + from;
}
static int label_index_drop(int pt /*1..5 = P,N,B,R,Q*/, int to) {
    static const int ch_of_pt[6] = {-1, 1, 2, 4, 3, 5}; /* label order P,N,R,B,Q */
    const int tr = to >> 3;
    if (pt < 1 || pt > 5) return -1;
    if (pt == 1 && (tr == 0 || tr == 7)) return -1;
    return ch_of_pt[pt] * 64 + to;
}

void hmo_policy_init(void) {
    if (g_policy_ready) return;
    for (int c = 0; c < 2; ++c) {
        const int m = c ? 56 : 0; /* globals.cc:107,115,126: mirror when color == 1 */
        for (int f = 0; f < 64; ++f)
            for (int t = 0; t < 64; ++t) {
                g_tab_normal[c][f][t][0] = label_index_normal(f ^ m, t ^ m);
                g_tab_normal[c][f][t][1] = label_index_knight_promo(f ^ m, t ^ m);
            }
        for (int t = 0; t < 64; ++t) {
            for (int pt = 0; pt < 8; ++pt) g_tab_drop[c][t][pt] = -1;
            for (int pt = 1; pt <= 5; ++pt) g_tab_drop[c][t][pt] = label_index_drop(pt, t ^ m);
        }
    }
    g_policy_ready = 1;
}

void hmo_policy_tables(int32_t* normal, int32_t* drop) {
    hmo_policy_init();
    memcpy(normal, g_tab_normal, sizeof(g_tab_normal));
    memcpy(drop, g_tab_drop, sizeof(g_tab_drop));
}

/* utils.h:184-216.  Move bits (Fairy-Stockfish/src/types.h:249-265,735-798 with
 * SQUARE_BITS = 6, MOVE_TYPE_BITS = 4): to [0:6], from [6:12], type [12:16]
 * (PROMOTION = 3, DROP = 4), piece type [16:22] (KNIGHT = 2, BISHOP = 3, ROOK = 4, QUEEN = 5). */
int32_t hmo_policy_index(uint32_t move, int stm_black) {
    hmo_policy_init();
    if (move == 0) return 0; /* MOVE_NONE -> pass */
    const int color = stm_black ? 1 : 0;
    const int to = (int)(move & 63u);           /* to_sq: m & SQUARE_BIT_MASK (63, types.h:489) */
    const int type = (int)((move >> 12) & 15u);
    const int pt = (int)((move >> 16) & 63u);
    if (type == 4) { /* DROP */
        if (to >= 0 && to < 64 && pt >= 1 && pt <= 5) return g_tab_drop[color][to][pt];
        return -1;
    }
    const int from = (int)((move >> 6) & 63u);  /* from_sq: (m >> SQUARE_BITS) & SQUARE_BIT_MASK */
    if (from >= 64 || to >= 64) return -1;
    if (type == 3) { /* PROMOTION */
        if (pt == 2) return g_tab_normal[color][from][to][1];
        if (pt == 5) return g_tab_normal[color][from][to][0];
        return -1;
    }
    return g_tab_normal[color][from][to][0];
}

/* utils.h:127-167 */
void hmo_normalize_logits(const float* logits, size_t n, float* out) {
    float max_logit = -INFINITY;
    for (size_t i = 0; i < n; ++i) {
        out[i] = 0.0f;
        if (isfinite(logits[i]) && logits[i] > max_logit) max_logit = logits[i];
    }
    if (!isfinite(max_logit)) {
        for (size_t i = 0; i < n; ++i) out[i] = 1.0f / (float)n;
        return;
    }
    double sum = 0.0;
    for (size_t i = 0; i < n; ++i)
        if (isfinite(logits[i])) {
            out[i] = expf(logits[i] - max_logit);
            sum += out[i];
        }
    if (!isfinite(sum) || sum <= 0.0) {
        size_t finite_count = 0;
        for (size_t i = 0; i < n; ++i) finite_count += isfinite(logits[i]) ? 1 : 0;
        if (finite_count > 0)
            for (size_t i = 0; i < n; ++i)
                out[i] = isfinite(logits[i]) ? 1.0f / (float)finite_count : 0.0f;
        return;
    }
    for (size_t i = 0; i < n; ++i) out[i] = (float)((double)out[i] / sum);
}

#define HMO_MAX_MOVES 1024
/* utils.h:226-243 */
void hmo_priors_f16(const uint16_t* policy, const uint32_t* moves, size_t n, int stm_black,
                    float* out) {
    float logits[HMO_MAX_MOVES];
    if (n > HMO_MAX_MOVES) n = HMO_MAX_MOVES;
    for (size_t i = 0; i < n; ++i) {
        int32_t idx = hmo_policy_index(moves[i], stm_black);
        logits[i] = idx >= 0 ? hmo_half_to_float(policy[idx]) : -INFINITY;
    }
    hmo_normalize_logits(logits, n, out);
}
void hmo_priors_f32(const float* policy, const uint32_t* moves, size_t n, int stm_black,
                    float* out) {
    float logits[HMO_MAX_MOVES];
    if (n > HMO_MAX_MOVES) n = HMO_MAX_MOVES;
    for (size_t i = 0; i < n; ++i) {
        int32_t idx = hmo_policy_index(moves[i], stm_black);
        logits[i] = idx >= 0 ? policy[idx] : -INFINITY;
    }
    hmo_normalize_logits(logits, n, out);
}

void hmo_batch_priors_f16(const uint16_t* logits_a, const uint16_t* logits_b,
                          const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                          const uint32_t* move_offsets, float* priors_out) {
    for (size_t i = 0; i < n_pos; ++i)
        for (int bn = 0; bn < 2; ++bn) {
            const uint32_t lo = move_offsets[2 * i + bn], hi = move_offsets[2 * i + bn + 1];
            const uint16_t* pol = (bn ? logits_b : logits_a) + i * HM_NB_POLICY_VALUES;
            const int stm_black = (pos[i].board[bn].flags & HM_FLAG_STM_BLACK) != 0;
            hmo_priors_f16(pol, moves + lo, hi - lo, stm_black, priors_out + lo);
        }
}

/* ------------------------------------------------------------------------------------
 * Value shaping — searchthread.cc:569-619
 * ---------------------------------------------------------------------------------- */
static float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

float hmo_shape_value(float value, const float* wdl, const float* moves_left,
                      const hm_value_config* cfg) {
    float scalar = isfinite(value) ? clampf(value, -1.0f, 1.0f) : 0.0f; /* :570-573 */
    float neural = scalar;
    if (cfg->enable_wdl_eval && wdl) { /* :576 */
        const float l = wdl[0], d = wdl[1], w = wdl[2];
        if (isfinite(l) && isfinite(d) && isfinite(w)) {
            float mx = l > d ? l : d;
            if (w > mx) mx = w;
            const float el = expf(l - mx), ed = expf(d - mx), ew = expf(w - mx);
            const float s = el + ed + ew;
            if (isfinite(s) && s > 0.0f) {
                const float pl = el / s, pd = ed / s, pw = ew / s;
                const float wdl_value = pw - pl - cfg->draw_contempt * pd;       /* :596-597 */
                const float wt = clampf(cfg->wdl_value_weight, 0.0f, 1.0f);      /* :599-600 */
                neural = (1.0f - wt) * scalar + wt * wdl_value;                   /* :602-604 */
            }
        }
    }
    if (moves_left && cfg->moves_left_discount > 0.0f) { /* :609 */
        const float plies = clampf(*moves_left, 0.0f, 1.0f);
        const float disc = clampf(cfg->moves_left_discount, 0.0f, 1.0f);
        neural *= 1.0f - disc * plies;
    }
    return clampf(neural, -1.0f, 1.0f); /* :619 */
}

float hmo_shape_value_half(uint16_t value, const uint16_t* wdl, const uint16_t* moves_left,
                           const hm_value_config* cfg) {
    float w[3], ml;
    if (wdl) { w[0] = hmo_half_to_float(wdl[0]); w[1] = hmo_half_to_float(wdl[1]); w[2] = hmo_half_to_float(wdl[2]); }
    if (moves_left) ml = hmo_half_to_float(*moves_left);
    return hmo_shape_value(hmo_half_to_float(value), wdl ? w : NULL, moves_left ? &ml : NULL, cfg);
}

/* ===================================================================================
 * Root policy stages (tools/selfplay.cc:167-185,301-376; search/node.h:288-315;
 * environment/joint_action.h:14-110)
 * =================================================================================== */
#include <stdlib.h>

int hmo_apply_temperature(float* p, size_t n, double temperature) {
    if (n == 0 || !(temperature > 0.0)) return 0;              /* selfplay.cc:168-170 */
    const double exponent = 1.0 / temperature;
    double total = 0.0;
    for (size_t i = 0; i < n; ++i) {                           /* :173-176 */
        p[i] = (float)pow((double)(p[i] > 0.0f ? p[i] : 0.0f), exponent);
        total += p[i];
    }
    if (!isfinite(total) || total <= 0.0) {                    /* :177-181 */
        const float uniform = 1.0f / (float)n;
        for (size_t i = 0; i < n; ++i) p[i] = uniform;
        return 1;
    }
    for (size_t i = 0; i < n; ++i) p[i] = (float)(p[i] / total); /* :182-184 */
    return 1;
}

void hmo_dirichlet_mix(float* p, const float* noise, size_t n, float epsilon) {
    if (n <= 1) return;                                        /* node.h:292-294 */
    volatile float total = 0.0f;                               /* volatile: one rounding per add */
    for (size_t i = 0; i < n; ++i) total = total + noise[i];   /* :299-303 */
    if (total <= 0.0f) return;                                 /* :304-306 */
    const float eps = epsilon < 0.0f ? 0.0f : (epsilon > 1.0f ? 1.0f : epsilon);
    for (size_t i = 0; i < n; ++i) {                           /* :308-311, op by op, no contraction */
        volatile float keep = (1.0f - eps) * p[i];
        volatile float scaled = eps * noise[i];
        volatile float add = scaled / total;
        p[i] = keep + add;
    }
}

size_t hmo_discrete_pick_f64(const double* w, size_t n, double u) {
    if (n < 2) return 0;                                       /* _M_prob.size() < 2: no table, no draw */
    double* cp = (double*)malloc(n * sizeof(double));
    double sum = 0.0;
    for (size_t i = 0; i < n; ++i) sum += w[i];                /* std::accumulate(.., 0.0) */
    double run = 0.0;
    for (size_t i = 0; i < n; ++i) { run += w[i] / sum; cp[i] = run; }  /* __normalize, partial_sum */
    cp[n - 1] = 1.0;
    size_t lo = 0, len = n;                                    /* std::lower_bound(cp, cp+n, u) */
    while (len > 0) {
        const size_t half = len >> 1;
        if (cp[lo + half] < u) { lo += half + 1; len -= half + 1; } else len = half;
    }
    free(cp);
    return lo;
}

size_t hmo_discrete_pick_f32(const float* p, size_t n, double u) {
    if (n < 2) return 0;
    double* w = (double*)malloc(n * sizeof(double));
    for (size_t i = 0; i < n; ++i) w[i] = (double)p[i];        /* vector<double>(first, last) */
    const size_t r = hmo_discrete_pick_f64(w, n, u);
    free(w);
    return r;
}

/* joint_action.h:14-32 */
static int double_sit_legal(int tadv, int a_on, int b_on) { return tadv && (a_on != b_on); }
static int single_pass_legal(int tadv, int a_on, int b_on, int partner_capture) {
    return tadv || !(a_on && b_on) || partner_capture;
}
/* joint_action.h:84-110: 1 when jointPrior >= 0 */
static int joint_is_legal(uint32_t m_a, uint32_t m_b, int cap_a, int cap_b, int tadv, int a_on, int b_on,
                          int a_can, int b_can) {
    const int sits_a = m_a == 0u, sits_b = m_b == 0u;
    if (sits_a && sits_b) return double_sit_legal(tadv, a_on, b_on);
    if (sits_a && a_can) return single_pass_legal(tadv, a_on, b_on, cap_b);
    if (sits_b && b_can) return single_pass_legal(tadv, a_on, b_on, cap_a);
    return 1;
}

int hmo_sample_joint(const hm_packed_pos* pos, const uint32_t* moves_a, const float* p_a, const uint8_t* cap_a,
                     size_t n_a, const uint32_t* moves_b, const float* p_b, const uint8_t* cap_b, size_t n_b,
                     const double* u, uint32_t* pick_a, uint32_t* pick_b) {
    *pick_a = *pick_b = 0;
    if (n_a == 0 || n_b == 0) return HM_ROOT_EMPTY_LIST;
    const uint8_t f_a = pos->board[0].flags, f_b = pos->board[1].flags;
    const int team_black = (f_a & HM_FLAG_TEAM_BLACK) != 0, tadv = (f_a & HM_FLAG_TIME_ADVANTAGE) != 0;
    const int a_on = ((f_a & HM_FLAG_STM_BLACK) != 0) == team_black;      /* selfplay.cc:310 */
    const int b_on = ((f_b & HM_FLAG_STM_BLACK) != 0) != team_black;      /* :311 */
    const int a_can = a_on && n_a > 1, b_can = b_on && n_b > 1;           /* :353-354 */
    for (int b = 0; b < 2; ++b) {
        const float* p = b ? p_b : p_a;
        const size_t n = b ? n_b : n_a;
        double s = 0.0;
        for (size_t i = 0; i < n; ++i) s += p[i];
        if (n >= 2 && (!(s > 0.0) || !isfinite(s))) return HM_ROOT_BAD_WEIGHTS;
    }
#define CAP(c, m, i) ((m)[i] != 0u && (c) != NULL && (c)[i] != 0)          /* :337-347 */
#define LEGAL(i, j) joint_is_legal(moves_a[i], moves_b[j], CAP(cap_a, moves_a, i), CAP(cap_b, moves_b, j), \
                                   tadv, a_on, b_on, a_can, b_can)
    size_t i_a = hmo_discrete_pick_f32(p_a, n_a, u[0]);                    /* :363-365 */
    size_t i_b = hmo_discrete_pick_f32(p_b, n_b, u[1]);
    int st = HM_ROOT_OK;
    if (!LEGAL(i_a, i_b)) {                                                /* :370-392 */
        size_t cap_pairs = n_a * n_b, n_legal = 0;
        size_t* pa = (size_t*)malloc(cap_pairs * sizeof(size_t));
        size_t* pb = (size_t*)malloc(cap_pairs * sizeof(size_t));
        double* w = (double*)malloc(cap_pairs * sizeof(double));
        double total = 0.0;
        for (size_t i = 0; i < n_a; ++i)
            for (size_t j = 0; j < n_b; ++j) {
                if (!LEGAL(i, j)) continue;
                pa[n_legal] = i; pb[n_legal] = j;
                w[n_legal] = (double)p_a[i] * (double)p_b[j];
                total += w[n_legal];
                ++n_legal;
            }
        if (!(total > 0.0) || !isfinite(total)) {
            st = HM_ROOT_NO_LEGAL_JOINT;
        } else {
            const size_t k = hmo_discrete_pick_f64(w, n_legal, u[2]);
            i_a = pa[k]; i_b = pb[k];
            st = n_legal >= 2 ? HM_ROOT_RESAMPLED : HM_ROOT_RESAMPLED_FORCED;
        }
        free(pa); free(pb); free(w);
    }
#undef LEGAL
#undef CAP
    *pick_a = (uint32_t)i_a; *pick_b = (uint32_t)i_b;
    return st;
}

/* MT19937-64 (Matsumoto & Nishimura) = std::mt19937_64; generate_canonical<double,53> takes ONE
 * 64-bit draw: double(x) / 2^64, and nextafter(1,0) if that rounds to 1 (bits/random.tcc). */
void hmo_mt19937_64_canonical(uint64_t seed, size_t n, double* out) {
    uint64_t mt[312];
    mt[0] = seed;
    for (int i = 1; i < 312; ++i) mt[i] = 6364136223846793005ULL * (mt[i - 1] ^ (mt[i - 1] >> 62)) + (uint64_t)i;
    int idx = 312;
    for (size_t k = 0; k < n; ++k) {
        if (idx >= 312) {
            for (int i = 0; i < 312; ++i) {
                const uint64_t x = (mt[i] & 0xFFFFFFFF80000000ULL) | (mt[(i + 1) % 312] & 0x7FFFFFFFULL);
                mt[i] = mt[(i + 156) % 312] ^ (x >> 1) ^ ((x & 1ULL) ? 0xB5026F5AA96619E9ULL : 0ULL);
            }
            idx = 0;
        }
        uint64_t x = mt[idx++];
        x ^= (x >> 29) & 0x5555555555555555ULL;
        x ^= (x << 17) & 0x71D67FFFEDA60000ULL;
        x ^= (x << 37) & 0xFFF7EEE000000000ULL;
        x ^= (x >> 43);
        double r = (double)x / 18446744073709551616.0;
        if (r >= 1.0) r = nextafter(1.0, 0.0);
        out[k] = r;
    }
}
