// render.cuh -- headless frame rasteriser: parallel_env.render, battle_env.py:498-560,
// into a 1200 x 800 RGB DEVICE buffer.  No window, no graphics API: the pygame display, the font
// (plane names, winner screen) and vidmaker are unavailable on this box and left out.
//
// Draw order of the reference: white background (:528), bullets (:531-532, 6x3 quads in the
// shooter's colour rotated by their heading, sprites.py:306-311,353-360), explosions (:535-536: the
// caller's list of {x, y, frame}, the animation history is host-side like self.explosions), both bases with their
// hp bar (:539-540, sprites.py:265-278), alive red then blue planes with their hp bar (:543-550,
// sprites.py:155-178).  Sprites are RGBA images handed over by the caller (the reference's
// assets/*.png or procedural stand-ins).  A plane sprite goes through pygame 2.1.2's transform.rotate algorithm
// (PlaneRot below: fixed-point inverse mapping, destination size rule, rotate90 for right angles) and is blitted at
// the rect blitRotate computes; bullets are 6 x 3 quads along their integer step (half-pixel coordinates, Q14 cos/sin:
// the env state keeps a bullet's integer step, not its direction).
//
// All geometry is integer, so the CPU restatement used by the tests reproduces every pixel bit for bit.  One thread per pixel; the
// env's objects are staged once per block in shared memory with their bounding boxes.
#pragma once
#include "tick.cuh"

namespace bs {

constexpr int R_W = 1200, R_H = 800;                 // sprites.py:9-10
constexpr int PLANE_W = 50, PLANE_H = 48, BASE_W = 62, BASE_H = 62, BUL_W = 6, BUL_H = 3;
constexpr int Q = 14;                                // fixed-point bits of cos / sin
constexpr int EXPL_W = 64, EXPL_FRAMES = 9, MAX_EXPL = 64;   // sprites.py:392-398

// pygame.transform.rotate of one plane sprite, prepared on the HOST by bs_render (sin / cos from the host libm, the
// functions pygame's C code calls): pygame 2.1.2 src_c/transform.c, surf_rotate() (destination size nx x ny from the
// rotated corner extents, truncated) and rotate() (16.16 fixed point: isin = (int)(sin * 65536), icos likewise, start
// ax = (nx << 15) - (int)(cos * ((nx - 1) << 15)), ay = (ny << 15) - (int)(sin * ((nx - 1) << 15)), recentring
// xd = (w - nx) << 15, yd = (h - ny) << 15; destination (x, y) samples source (dx >> 16, dy >> 16) with
// dx = ax + xd + isin * (ny / 2 - y) + icos * x, dy = ay + yd - icos * (ny / 2 - y) + isin * x, transparent outside);
// multiples of 90 degrees take rotate90()'s exact pixel permutation (turns = 0..3, else -1).  The rotated surface is
// blitted at the rect blitRotate computes (sprites.py:13-33): centre = the plane's rect centre, i.e. left = cx - (nx >> 1).
struct PlaneRot { int left, top, nx, ny, ax, ay, isin, icos, turns; };

struct RenderArgs {
    long long env;
    PlaneRot rot[16];              // per plane of the env
    const uint8_t *plane_img[2];   // [48][50][4] RGBA, red / blue
    const uint8_t *base_img[2];    // [62][62][4]
    uint8_t *rgb;                  // [800][1200][3]
    const uint8_t *expl_img;       // [9][64][64][4] or nullptr
    const int32_t *expl;           // [n_expl][3] = x, y, frame
    int n_expl;
};

struct RObj {            // one rotated rectangle: plane or bullet
    int cx2, cy2;        // centre in half-pixels (2 * rect.center)
    int c, s;            // Q14 cos / sin of the heading
    int w, h;            // sprite size
    int rad2;            // bounding radius in half-pixels
    int kind;            // 0/1: plane sprite of team; 2/3: bullet of team
};

struct RBar { int x0, y0, x1, y1, team; };  // filled rect [x0,x1) x [y0,y1) with a 1-px black border around it

__device__ __forceinline__ void blend(uint8_t (&px)[3], const uint8_t *src) {
    // pygame per-pixel-alpha blit onto an opaque surface: d += ((s - d) * a + s) >> 8
    const int a = src[3];
    if (a == 0) return;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int d = px[c], s = src[c];
        px[c] = (uint8_t)(d + ((((s - d) * a) + s) >> 8));
    }
}

__device__ __forceinline__ void team_colour(int team, uint8_t (&px)[3]) {  // sprites.py:6-7 RED / BLUE
    px[0] = team == 0 ? 138 : 0; px[1] = team == 0 ? 24 : 93; px[2] = team == 0 ? 26 : 135;
}

// hp bar -- sprites.py:171-178, 269-275: rect hp*10 x 10 centred `dy` above the sprite centre inside a
// black (hp*10+2) x 12 rect (border_radius 3 is not reproduced: square corners).
__device__ __forceinline__ RBar make_bar(int cx, int cy, int dy, int hp, int team) {
    RBar b;
    const int w = hp * 10, h = 10, ccx = cx, ccy = cy - dy;
    b.x0 = ccx - (w >> 1); b.y0 = ccy - (h >> 1); b.x1 = b.x0 + w; b.y1 = b.y0 + h; b.team = hp > 0 ? team : -1;
    return b;
}

template <int N>
__global__ void __launch_bounds__(256) render_kernel(Params p, RenderArgs ra) {
    constexpr int P = 2 * N, MAXO = P + P * SLOTS;
    __shared__ RObj obj[MAXO];
    __shared__ RBar bar[P + 2];
    __shared__ int base_xy[2][2];
    __shared__ int expl[MAX_EXPL][3];
    const long long e = ra.env;
    const int n_expl = ra.expl_img != nullptr && ra.expl != nullptr ? min(ra.n_expl, MAX_EXPL) : 0;
    for (int q = threadIdx.x; q < n_expl * 3; q += blockDim.x) expl[q / 3][q % 3] = ra.expl[q];
    // ---- stage the env's objects: bullets first (drawn first), then planes -----------------------
    for (int q = threadIdx.x; q < P * SLOTS; q += blockDim.x) obj[q].kind = -1;
    __syncthreads();
    {
        // the env's bullets are the entries of its warp's pool whose owner lane belongs to it; they are
        // staged at [age index][agent] -- list order of self.bullets is spawn order, and any fixed order
        // gives the same picture unless two bullets of different teams overlap
        constexpr int K = 32 / P;
        const long long gw = e / K;
        const int lane0 = (int)(e - gw * K) * P;
        const int cnt = p.pool_cnt[gw];
        for (int q = threadIdx.x; q < cnt; q += blockDim.x) {
            const uint2 en = p.pool[gw * POOL_CAP + q];
            const uint32_t Pw = en.x, M = en.y;
            const int owner = (M >> 23) & 31, age = (int)(M >> 28);
            if (!(Pw & B_VALID) || owner < lane0 || owner >= lane0 + P || age < 1 || age > SLOTS) continue;
            const int i = owner - lane0, ai = SLOTS - age;
            const int32_t D = meta_dlt(M);
            int sx = (int16_t)(D & 0xffff), sy = (D - sx) >> 16;
            const uint32_t Qw = (Pw & B_POS) - B_BIAS;
            const double len = sqrt((double)(sx * sx + sy * sy));
            RObj o;
            o.cx2 = 2 * (int)(Qw & 0xffffu); o.cy2 = 2 * (int)(Qw >> 16);
            // heading of the integer step: screen y grows downwards, angles counter-clockwise
            o.c = len > 0 ? (int)rint((double)sx / len * (1 << Q)) : (1 << Q);
            o.s = len > 0 ? (int)rint((double)-sy / len * (1 << Q)) : 0;
            o.w = BUL_W; o.h = BUL_H; o.rad2 = 8; o.kind = 2 + (i >= N);
            obj[ai * P + i] = o;
        }
    }
    for (int i = threadIdx.x; i < P; i += blockDim.x) {
        const uint32_t pos = p.plane_pos[e * P + i];
        const int hp = p.plane_hp[e * P + i];
        const double dir = p.plane_dir[e * P + i];
        RObj o;
        o.cx2 = 2 * (int)(pos & 0xffff); o.cy2 = 2 * (int)(pos >> 16);
        o.c = (int)rint(cos(dir * DEG2RAD) * (1 << Q)); o.s = (int)rint(sin(dir * DEG2RAD) * (1 << Q));
        o.w = PLANE_W; o.h = PLANE_H; o.rad2 = 72; o.kind = hp > 0 ? (i >= N) : -1;   // only alive planes, :543-550
        obj[P * SLOTS + i] = o;
        bar[2 + i] = make_bar((int)(pos & 0xffff), (int)(pos >> 16), 35, hp > 0 ? hp : 0, i >= N);
    }
    if (threadIdx.x < 2) {
        const int t = threadIdx.x;
        const uint32_t bp = p.base_pos[e * 2 + t];
        base_xy[t][0] = bp & 0xffff; base_xy[t][1] = bp >> 16;
        const int hp = p.base_hp[e * 2 + t];
        bar[t] = make_bar(bp & 0xffff, bp >> 16, 40, hp > 0 ? hp : 0, t);
    }
    __syncthreads();

    const int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= R_W * R_H) return;
    const int py = pix / R_W, px_ = pix - py * R_W;
    uint8_t out[3] = {255, 255, 255};  // sprites.WHITE

    auto rotated = [&](const RObj &o) {   // bullets: 6 x 3 quads along their integer step (the state keeps no bullet direction)
        if (o.kind < 0) return;
        const int dx = 2 * px_ + 1 - o.cx2, dy = 2 * py + 1 - o.cy2;   // pixel centre - sprite centre, half-pixels
        if (abs(dx) > o.rad2 || abs(dy) > o.rad2) return;
        // inverse of a counter-clockwise rotation by the heading on a y-down screen
        const int ux = dx * o.c - dy * o.s, uy = dx * o.s + dy * o.c;       // Q14 half-pixels
        const int u = (ux + (o.w << Q)) >> (Q + 1), v = (uy + (o.h << Q)) >> (Q + 1);  // floor((u/2) + w/2)
        if (u < 0 || u >= o.w || v < 0 || v >= o.h) return;
        team_colour(o.kind - 2, out);
    };
    auto plane = [&](const RObj &o, const PlaneRot &r) {   // pygame.transform.rotate + blit, see PlaneRot
        if (o.kind < 0) return;
        const int x = px_ - r.left, y = py - r.top;
        if (x < 0 || x >= r.nx || y < 0 || y >= r.ny) return;
        int u, v;
        if (r.turns >= 0) {                                 // rotate90(): dst(x, y) = src(u, v)
            u = r.turns == 0 ? x : (r.turns == 1 ? PLANE_W - 1 - y : (r.turns == 2 ? PLANE_W - 1 - x : y));
            v = r.turns == 0 ? y : (r.turns == 1 ? x : (r.turns == 2 ? PLANE_H - 1 - y : PLANE_H - 1 - x));
        } else {
            const int cyd = r.ny / 2;
            const int sx = r.ax + r.isin * (cyd - y) + r.icos * x, sy = r.ay - r.icos * (cyd - y) + r.isin * x;
            if (sx < 0 || sy < 0 || sx > (PLANE_W << 16) - 1 || sy > (PLANE_H << 16) - 1) return;   // bgcolor: transparent
            u = sx >> 16; v = sy >> 16;
        }
        blend(out, ra.plane_img[o.kind] + (v * PLANE_W + u) * 4);
    };
    auto draw_bar = [&](const RBar &b) {
        if (b.team < 0) return;
        if (px_ >= b.x0 - 1 && px_ < b.x1 + 1 && py >= b.y0 - 1 && py < b.y1 + 1) {
            if (px_ >= b.x0 && px_ < b.x1 && py >= b.y0 && py < b.y1) team_colour(b.team, out);
            else { out[0] = out[1] = out[2] = 0; }
        }
    };
    for (int q = 0; q < P * SLOTS; ++q) rotated(obj[q]);
    for (int q = 0; q < n_expl; ++q) {  // Explosion.draw, sprites.py:405-419: frame blitted at rect centred on the kill
        const int f = expl[q][2], u = px_ - (expl[q][0] - EXPL_W / 2), v = py - (expl[q][1] - EXPL_W / 2);
        if (f >= 0 && f < EXPL_FRAMES && u >= 0 && u < EXPL_W && v >= 0 && v < EXPL_W)
            blend(out, ra.expl_img + ((size_t)(f * EXPL_W + v) * EXPL_W + u) * 4);
    }
#pragma unroll
    for (int t = 0; t < 2; ++t) {  // Base.draw, sprites.py:265-275: blit at rect (centre - 31), then the bar
        const int u = px_ - (base_xy[t][0] - BASE_W / 2), v = py - (base_xy[t][1] - BASE_H / 2);
        if (u >= 0 && u < BASE_W && v >= 0 && v < BASE_H) blend(out, ra.base_img[t] + (v * BASE_W + u) * 4);
        draw_bar(bar[t]);
    }
    for (int i = 0; i < P; ++i) { plane(obj[P * SLOTS + i], ra.rot[i]); if (obj[P * SLOTS + i].kind >= 0) draw_bar(bar[2 + i]); }
    uint8_t *dst = ra.rgb + (size_t)pix * 3;
    dst[0] = out[0]; dst[1] = out[1]; dst[2] = out[2];
}

}  // namespace bs
