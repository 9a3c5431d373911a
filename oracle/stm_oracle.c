/*
 * stm_oracle.c — CPU ORACLE.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product path (softtreemax_b200/) never does.
 *
 * What it restates
 * ----------------
 * The tree-expansion hot path of NVlabs/SoftTreeMax:
 *   - tree semantics: policies/cule_bfs.py:98-209 (root injection :110-118, child/parent/action
 *     order :90,141-144, frameskip loop with reward weight gamma**depth :152-158, frames taken at
 *     skip index frameskip-2 / frameskip-1 :162-167, max :168, stack shift :170-172, sign of the
 *     cumulative reward :204-206);
 *   - the emulator those lines drive.  That arithmetic lives in a third-party dependency that is
 *     NOT under /root/reference and is not installable here: NVLabs/cule, branch `bfs`, no pinned
 *     commit (reference Dockerfile:7-8), Python package `torchcule`.  The reference holds no tests,
 *     golden vectors or fixtures.
 *
 * PARITY UNPINNED for the emulator (RAM / frames / rewards / done): there is nothing of the
 * reference's to pin it against on this box.  The emulator below is a restatement of the public
 * Atari 2600 hardware behaviour (6507 ISA + cycle counts, RIOT 6532, TIA) under the explicit,
 * simplified timing model written down in DESIGN.md §3; the CUDA path is required to match THIS
 * model bit for bit.  The oracle itself is pinned by self-authored known-answer tests
 * (tests/test_oracle_*.py) and by the golden trace hashes in tests/golden/.
 *
 * Style: one switch per opcode, per-pixel TIA evaluation — deliberately the textbook structure,
 * so that it is an independent check on the CUDA implementation (which decodes by addressing
 * mode / operation class and renders with 160-bit object masks).
 */
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
#include <math.h>
#include "../include/stm_state.h"
#include "../include/stm_palette.h"

#ifdef _OPENMP
#include <omp.h>
#endif

typedef stm_env_state env_t;

typedef struct orc_game {
    stm_game_desc d;
    uint8_t rom[8192];
    uint16_t rom_mask;   /* address mask inside one 4K (or 2K) window */
} orc_game;

static uint8_t GRAY[128] = STM_GRAY_LUT_INIT;
/* the grey table is data (stm_set_palette on the CUDA side): tests install the same table on both sides */
void orc_set_palette(const uint8_t *grey128) { memcpy(GRAY, grey128, 128); }

/* status flags */
enum { FC = 1, FZ = 2, FI = 4, FD = 8, FB = 16, FU = 32, FV = 64, FN = 128 };

typedef struct cpu_ctx {
    const orc_game *g;
    env_t *e;
    uint8_t *fb;         /* native frame being rendered (160x210 colour indices) or NULL */
    int wsync;           /* the current instruction wrote WSYNC */
    int end_line_no;     /* scanline number at which the frame ended (set with STM_F_FRAME_DONE) */
} cpu_ctx;

size_t orc_game_size(void) { return sizeof(orc_game); }

int orc_game_init(orc_game *g, const stm_game_desc *d, const uint8_t *rom)
{
    memset(g, 0, sizeof(*g));
    g->d = *d;
    if (d->bank_scheme == STM_BANK_NONE) {
        if (d->rom_size != 2048 && d->rom_size != 4096) return -1;
        g->rom_mask = (uint16_t)(d->rom_size - 1);
    } else if (d->bank_scheme == STM_BANK_F8) {
        if (d->rom_size != 8192) return -1;
        g->rom_mask = 0x0FFF;
    } else return -1;
    if (d->n_actions < 1 || d->n_actions > 18) return -1;
    memcpy(g->rom, rom, d->rom_size);
    return 0;
}

/* ------------------------------------------------------------------------------------------- */
/* TIA: per-pixel evaluation                                                                    */
/* ------------------------------------------------------------------------------------------- */

static const uint8_t COPY_N[8]      = {1, 2, 2, 3, 2, 1, 3, 1};
static const uint8_t COPY_OFF[8][3] = {{0,0,0},{0,16,0},{0,32,0},{0,16,32},{0,64,0},{0,0,0},{0,32,64},{0,0,0}};

static int pf_at(const env_t *e, int x)
{
    int i = x < 80 ? (x >> 2) : ((e->ctrlpf & 1) ? 19 - ((x - 80) >> 2) : ((x - 80) >> 2));
    if (i < 4)  return (e->pf0 >> (4 + i)) & 1;
    if (i < 12) return (e->pf1 >> (11 - i)) & 1;
    return (e->pf2 >> (i - 12)) & 1;
}

static int player_at(int x, int pos, uint8_t grp, uint8_t nusiz, uint8_t refp)
{
    int mode = nusiz & 7;
    int scale = mode == 5 ? 2 : (mode == 7 ? 4 : 1);
    if (!grp) return 0;
    for (int c = 0; c < COPY_N[mode]; c++) {
        int d = (x - pos - COPY_OFF[mode][c] + 320) % 160;
        if (d < 8 * scale) {
            int b = d / scale;
            int bit = (refp & 8) ? b : 7 - b;
            if ((grp >> bit) & 1) return 1;
        }
    }
    return 0;
}

static int missile_at(int x, int pos, uint8_t enam, uint8_t nusiz, uint8_t resmp)
{
    if (!(enam & 2) || (resmp & 2)) return 0;
    int mode = nusiz & 7;
    int w = 1 << ((nusiz >> 4) & 3);
    for (int c = 0; c < COPY_N[mode]; c++) {
        int d = (x - pos - COPY_OFF[mode][c] + 320) % 160;
        if (d < w) return 1;
    }
    return 0;
}

static int ball_at(const env_t *e, int x)
{
    uint8_t en = (e->vdelbl & 1) ? e->enabl_old : e->enabl;
    if (!(en & 2)) return 0;
    int w = 1 << ((e->ctrlpf >> 4) & 3);
    return ((x - e->pos[STM_BL] + 160) % 160) < w;
}

/* render pixels [e->rx, xe) of the current line with the registers as they are now */
static void tia_catchup(cpu_ctx *c, int xe)
{
    env_t *e = c->e;
    if (xe > STM_VIS_W) xe = STM_VIS_W;
    if (xe <= e->rx) return;
    int row = (int)e->line - STM_YSTART;
    uint8_t *dst = (c->fb && row >= 0 && row < STM_VIS_H) ? c->fb + row * STM_VIS_W : NULL;
    uint8_t g0 = (e->vdelp0 & 1) ? e->grp0_old : e->grp0;
    uint8_t g1 = (e->vdelp1 & 1) ? e->grp1_old : e->grp1;
    int vb = e->vblank & 2;
    if (!dst) {
        /* nothing to draw: collisions are the only effect, and they need VBLANK off and at least
           one movable object switched on */
        uint8_t en = (e->vdelbl & 1) ? e->enabl_old : e->enabl;
        if (vb || !(g0 | g1 | ((e->enam0 | e->enam1 | en) & 2))) { e->rx = (uint8_t)xe; return; }
    }
    for (int x = e->rx; x < xe; x++) {
        int pf = pf_at(e, x);
        int bl = ball_at(e, x);
        int p0 = player_at(x, e->pos[STM_P0], g0, e->nusiz0, e->refp0);
        int p1 = player_at(x, e->pos[STM_P1], g1, e->nusiz1, e->refp1);
        int m0 = missile_at(x, e->pos[STM_M0], e->enam0, e->nusiz0, e->resmp0);
        int m1 = missile_at(x, e->pos[STM_M1], e->enam1, e->nusiz1, e->resmp1);
        if (!vb) {
            uint16_t cx = 0;
            if (m0 & p0) cx |= 1 << 0;   /* CXM0P  D6 */
            if (m0 & p1) cx |= 1 << 1;   /* CXM0P  D7 */
            if (m1 & p1) cx |= 1 << 2;   /* CXM1P  D6 */
            if (m1 & p0) cx |= 1 << 3;   /* CXM1P  D7 */
            if (p0 & bl) cx |= 1 << 4;   /* CXP0FB D6 */
            if (p0 & pf) cx |= 1 << 5;   /* CXP0FB D7 */
            if (p1 & bl) cx |= 1 << 6;   /* CXP1FB D6 */
            if (p1 & pf) cx |= 1 << 7;   /* CXP1FB D7 */
            if (m0 & bl) cx |= 1 << 8;   /* CXM0FB D6 */
            if (m0 & pf) cx |= 1 << 9;   /* CXM0FB D7 */
            if (m1 & bl) cx |= 1 << 10;  /* CXM1FB D6 */
            if (m1 & pf) cx |= 1 << 11;  /* CXM1FB D7 */
            if (bl & pf) cx |= 1 << 13;  /* CXBLPF D7 */
            if (m0 & m1) cx |= 1 << 14;  /* CXPPMM D6 */
            if (p0 & p1) cx |= 1 << 15;  /* CXPPMM D7 */
            e->cx |= cx;
        }
        if (dst) {
            uint8_t col;
            if (vb) col = 0;
            else if ((e->flags & STM_F_HMOVE_BLANK) && x < 8) col = 0;
            else if (e->ctrlpf & 4) {              /* playfield/ball priority */
                if (pf | bl) col = e->colupf;
                else if (p0 | m0) col = e->colup0;
                else if (p1 | m1) col = e->colup1;
                else col = e->colubk;
            } else {
                if (p0 | m0) col = e->colup0;
                else if (p1 | m1) col = e->colup1;
                else if (bl) col = e->colupf;
                else if (pf) col = (e->ctrlpf & 2) ? (x < 80 ? e->colup0 : e->colup1) : e->colupf;
                else col = e->colubk;
            }
            dst[x] = col >> 1;
        }
    }
    e->rx = (uint8_t)xe;
}

static void end_line(cpu_ctx *c)
{
    env_t *e = c->e;
    tia_catchup(c, STM_VIS_W);
    e->line++;
    e->rx = 0;
    e->flags &= (uint8_t)~STM_F_HMOVE_BLANK;
    if (!(e->vblank & 0x80) && e->charge < 255) e->charge++;
    if (e->line >= STM_MAX_LINES) { e->flags |= STM_F_FRAME_DONE; c->end_line_no = STM_MAX_LINES; e->line = 0; }
}

static void advance(cpu_ctx *c, int cyc)
{
    env_t *e = c->e;
    e->timer -= cyc;
    if (e->timer < 0) e->tflag |= 0x80;
    if (e->timer < -(1 << 30)) e->timer += (1 << 30);   /* keeps (timer & 0xFF); avoids int32 wrap */
    int h = e->hcyc + cyc;
    while (h >= STM_CYC_PER_LINE) { end_line(c); h -= STM_CYC_PER_LINE; }
    e->hcyc = (uint8_t)h;
}

static int cur_x(const env_t *e)
{
    int x = e->hcyc * 3 - STM_HBLANK_CLK;
    return x < 0 ? 0 : x;
}

static int fire_pressed(const env_t *e)
{
    uint8_t a = e->action;
    return a == 1 || a >= 10;
}
static void action_dirs(uint8_t a, int *up, int *down, int *left, int *right)
{
    /* ALE action codes: 0 NOOP 1 FIRE 2 UP 3 RIGHT 4 LEFT 5 DOWN 6 UPRIGHT 7 UPLEFT 8 DOWNRIGHT
       9 DOWNLEFT 10 UPFIRE 11 RIGHTFIRE 12 LEFTFIRE 13 DOWNFIRE 14..17 diagonals+FIRE */
    static const uint8_t D[18] = { 0, 0, 1, 8, 4, 2, 9, 5, 10, 6, 1, 8, 4, 2, 9, 5, 10, 6 };
    uint8_t m = a < 18 ? D[a] : 0;       /* bit0 up, bit1 down, bit2 left, bit3 right */
    *up = m & 1; *down = (m >> 1) & 1; *left = (m >> 2) & 1; *right = (m >> 3) & 1;
}

static void update_latch(env_t *e, const orc_game *g)
{
    if (!(e->vblank & 0x40)) { e->inpt_latch = 0; return; }
    if (g->d.controller == STM_CTRL_JOYSTICK && fire_pressed(e)) e->inpt_latch |= 1;
}

static uint8_t tia_read(cpu_ctx *c, int reg)
{
    env_t *e = c->e;
    reg &= 0x0F;
    if (reg < 8) {
        tia_catchup(c, cur_x(e));
        return (uint8_t)(((e->cx >> (2 * reg)) & 3) << 6);
    }
    if (reg < 12) {                          /* INPT0..3: paddles */
        int i = reg - 8;
        if (c->g->d.controller != STM_CTRL_PADDLE || i > 1) return 0x00;
        if (e->vblank & 0x80) return 0x00;
        return e->charge >= e->paddle[i] ? 0x80 : 0x00;
    }
    if (reg == 12) {                         /* INPT4: joystick 0 fire */
        if (c->g->d.controller != STM_CTRL_JOYSTICK) return 0x80;
        if (fire_pressed(e)) return 0x00;
        if ((e->vblank & 0x40) && (e->inpt_latch & 1)) return 0x00;
        return 0x80;
    }
    if (reg == 13) return 0x80;
    return 0x00;
}

static void set_pos(env_t *e, int obj, int is_player)
{
    int clk = e->hcyc * 3;
    if (clk < STM_HBLANK_CLK) e->pos[obj] = is_player ? 3 : 2;
    else e->pos[obj] = (uint8_t)((clk - STM_HBLANK_CLK + (is_player ? 5 : 4)) % 160);
}

static void hmove_obj(env_t *e, int obj, uint8_t hm)
{
    int m = (int8_t)hm >> 4;                  /* +7 .. -8, positive = left */
    e->pos[obj] = (uint8_t)((e->pos[obj] - m + 160) % 160);
}

static void tia_write(cpu_ctx *c, int reg, uint8_t v)
{
    env_t *e = c->e;
    reg &= 0x3F;
    if (reg > 0x2C) return;
    if (reg == 0x02) return;                 /* WSYNC handled by the caller after the instruction */
    if (reg == 0x03 || (reg >= 0x15 && reg <= 0x1A)) return;   /* RSYNC, audio: ignored */
    tia_catchup(c, cur_x(e));
    switch (reg) {
    case 0x00:
        if ((v & 2) && !(e->vsync & 2)) { e->flags |= STM_F_FRAME_DONE; c->end_line_no = e->line; e->line = 0; }
        e->vsync = v; break;
    case 0x01:
        if ((v & 0x80) ) e->charge = 0;
        e->vblank = v; update_latch(e, c->g); break;
    case 0x04: e->nusiz0 = v; break;
    case 0x05: e->nusiz1 = v; break;
    case 0x06: e->colup0 = v; break;
    case 0x07: e->colup1 = v; break;
    case 0x08: e->colupf = v; break;
    case 0x09: e->colubk = v; break;
    case 0x0A: e->ctrlpf = v; break;
    case 0x0B: e->refp0 = v; break;
    case 0x0C: e->refp1 = v; break;
    case 0x0D: e->pf0 = v; break;
    case 0x0E: e->pf1 = v; break;
    case 0x0F: e->pf2 = v; break;
    case 0x10: set_pos(e, STM_P0, 1); break;
    case 0x11: set_pos(e, STM_P1, 1); break;
    case 0x12: set_pos(e, STM_M0, 0); break;
    case 0x13: set_pos(e, STM_M1, 0); break;
    case 0x14: set_pos(e, STM_BL, 0); break;
    case 0x1B: e->grp0 = v; e->grp1_old = e->grp1; break;
    case 0x1C: e->grp1 = v; e->grp0_old = e->grp0; e->enabl_old = e->enabl; break;
    case 0x1D: e->enam0 = v; break;
    case 0x1E: e->enam1 = v; break;
    case 0x1F: e->enabl = v; break;
    case 0x20: e->hmp0 = v; break;
    case 0x21: e->hmp1 = v; break;
    case 0x22: e->hmm0 = v; break;
    case 0x23: e->hmm1 = v; break;
    case 0x24: e->hmbl = v; break;
    case 0x25: e->vdelp0 = v; break;
    case 0x26: e->vdelp1 = v; break;
    case 0x27: e->vdelbl = v; break;
    case 0x28:
        if ((e->resmp0 & 2) && !(v & 2)) e->pos[STM_M0] = (uint8_t)((e->pos[STM_P0] + 4) % 160);
        e->resmp0 = v; break;
    case 0x29:
        if ((e->resmp1 & 2) && !(v & 2)) e->pos[STM_M1] = (uint8_t)((e->pos[STM_P1] + 4) % 160);
        e->resmp1 = v; break;
    case 0x2A:
        hmove_obj(e, STM_P0, e->hmp0); hmove_obj(e, STM_P1, e->hmp1);
        hmove_obj(e, STM_M0, e->hmm0); hmove_obj(e, STM_M1, e->hmm1);
        hmove_obj(e, STM_BL, e->hmbl);
        if (e->hcyc * 3 < STM_HBLANK_CLK) e->flags |= STM_F_HMOVE_BLANK;
        break;
    case 0x2B: e->hmp0 = e->hmp1 = e->hmm0 = e->hmm1 = e->hmbl = 0; break;
    case 0x2C: e->cx = 0; break;
    default: break;
    }
}

/* ------------------------------------------------------------------------------------------- */
/* RIOT                                                                                         */
/* ------------------------------------------------------------------------------------------- */
static uint8_t swcha_in(const cpu_ctx *c)
{
    const env_t *e = c->e;
    if (c->g->d.controller == STM_CTRL_PADDLE)
        return fire_pressed(e) ? 0x7F : 0xFF;
    int up, down, left, right;
    action_dirs(e->action, &up, &down, &left, &right);
    uint8_t v = 0xFF;
    if (right) v &= (uint8_t)~0x80;
    if (left)  v &= (uint8_t)~0x40;
    if (down)  v &= (uint8_t)~0x20;
    if (up)    v &= (uint8_t)~0x10;
    return v;
}

static uint8_t riot_read(cpu_ctx *c, uint16_t addr)
{
    env_t *e = c->e;
    if (addr & 0x04) {
        if (addr & 1) return e->tflag;                     /* INSTAT */
        uint8_t v = e->timer >= 0 ? (uint8_t)(e->timer >> e->tshift) : (uint8_t)(e->timer & 0xFF);
        e->tflag &= 0x7F;                                   /* reading INTIM clears the latch */
        return v;
    }
    switch (addr & 3) {
    case 0: return (uint8_t)((e->swcha_out & e->swacnt) | (swcha_in(c) & ~e->swacnt));
    case 1: return e->swacnt;
    case 2: return (uint8_t)((e->swchb_out & e->swbcnt) | (0x0B & ~e->swbcnt));
    default: return e->swbcnt;
    }
}

static void riot_write(cpu_ctx *c, uint16_t addr, uint8_t v)
{
    env_t *e = c->e;
    if (addr & 0x04) {
        if (addr & 0x10) {
            static const uint8_t SH[4] = {0, 3, 6, 10};
            e->tshift = SH[addr & 3];
            e->timer = ((int32_t)v << e->tshift) - 1;
            e->tflag &= 0x7F;
        }
        return;
    }
    switch (addr & 3) {
    case 0: e->swcha_out = v; break;
    case 1: e->swacnt = v; break;
    case 2: e->swchb_out = v; break;
    default: e->swbcnt = v; break;
    }
}

/* ------------------------------------------------------------------------------------------- */
/* bus                                                                                          */
/* ------------------------------------------------------------------------------------------- */
static uint8_t rom_at(const orc_game *g, const env_t *e, uint16_t addr)
{
    return g->rom[((uint32_t)e->bank << 12) | (addr & g->rom_mask)];
}

static void hotspot(cpu_ctx *c, uint16_t addr)
{
    if (c->g->d.bank_scheme == STM_BANK_F8) {
        uint16_t a = addr & 0x1FFF;
        if (a == 0x1FF8) c->e->bank = 0;
        else if (a == 0x1FF9) c->e->bank = 1;
    }
}

/* side-effect-free read used for instruction fetch and pointer fetch */
static uint8_t peek(cpu_ctx *c, uint16_t addr)
{
    addr &= 0x1FFF;
    if (addr & 0x1000) return rom_at(c->g, c->e, addr);
    if ((addr & 0x80) && !(addr & 0x200)) return c->e->ram[addr & 0x7F];
    return 0;
}

static uint8_t rd(cpu_ctx *c, uint16_t addr)
{
    addr &= 0x1FFF;
    if (addr & 0x1000) { hotspot(c, addr); return rom_at(c->g, c->e, addr); }
    if (!(addr & 0x80)) return tia_read(c, addr);
    if (!(addr & 0x200)) return c->e->ram[addr & 0x7F];
    return riot_read(c, addr);
}

static void wr(cpu_ctx *c, uint16_t addr, uint8_t v)
{
    addr &= 0x1FFF;
    if (addr & 0x1000) { hotspot(c, addr); return; }
    if (!(addr & 0x80)) { if ((addr & 0x3F) == 0x02) c->wsync = 1; tia_write(c, addr, v); return; }
    if (!(addr & 0x200)) { c->e->ram[addr & 0x7F] = v; return; }
    riot_write(c, addr, v);
}

/* ------------------------------------------------------------------------------------------- */
/* 6507                                                                                         */
/* ------------------------------------------------------------------------------------------- */
#define SETNZ(v) do { e->p = (uint8_t)((e->p & ~(FN | FZ)) | ((v) & 0x80) | (((v) & 0xFF) ? 0 : FZ)); } while (0)

static void op_adc(env_t *e, uint8_t v)
{
    unsigned a = e->a, cin = e->p & FC;
    if (e->p & FD) {
        unsigned al = (a & 0x0F) + (v & 0x0F) + cin;
        if (al > 9) al += 6;
        unsigned ah = (a >> 4) + (v >> 4) + (al > 0x0F);
        unsigned bin = (a + v + cin) & 0xFF;
        e->p &= (uint8_t)~(FN | FZ | FV | FC);
        if (!bin) e->p |= FZ;
        if (ah & 8) e->p |= FN;
        if (~(a ^ v) & (a ^ (ah << 4)) & 0x80) e->p |= FV;
        if (ah > 9) ah += 6;
        if (ah > 15) e->p |= FC;
        e->a = (uint8_t)((ah << 4) | (al & 0x0F));
    } else {
        unsigned s = a + v + cin;
        e->p &= (uint8_t)~(FV | FC);
        if (s > 0xFF) e->p |= FC;
        if (~(a ^ v) & (a ^ s) & 0x80) e->p |= FV;
        e->a = (uint8_t)s;
        SETNZ(e->a);
    }
}

static void op_sbc(env_t *e, uint8_t v)
{
    unsigned a = e->a, bin_c = (e->p & FC) ? 0 : 1;
    unsigned s = a - v - bin_c;
    uint8_t res = (uint8_t)s;
    int dec = e->p & FD;
    e->p &= (uint8_t)~(FV | FC);
    if (!(s & 0x100)) e->p |= FC;
    if ((a ^ v) & (a ^ s) & 0x80) e->p |= FV;
    SETNZ(res);
    if (dec) {
        int al = (int)(a & 0x0F) - (int)(v & 0x0F) - (int)bin_c;
        int ah = (int)(a >> 4) - (int)(v >> 4);
        if (al & 0x10) { al -= 6; ah--; }
        if (ah & 0x10) ah -= 6;
        res = (uint8_t)(((ah << 4) | (al & 0x0F)) & 0xFF);
    }
    e->a = res;
}

static void op_cmp(env_t *e, uint8_t r, uint8_t v)
{
    unsigned d = (unsigned)r - v;
    e->p &= (uint8_t)~FC;
    if (!(d & 0x100)) e->p |= FC;
    SETNZ(d);
}

static uint8_t op_asl(env_t *e, uint8_t v) { e->p = (uint8_t)((e->p & ~FC) | (v >> 7)); v <<= 1; SETNZ(v); return v; }
static uint8_t op_lsr(env_t *e, uint8_t v) { e->p = (uint8_t)((e->p & ~FC) | (v & 1)); v >>= 1; SETNZ(v); return v; }
static uint8_t op_rol(env_t *e, uint8_t v) { unsigned c = e->p & FC; e->p = (uint8_t)((e->p & ~FC) | (v >> 7)); v = (uint8_t)((v << 1) | c); SETNZ(v); return v; }
static uint8_t op_ror(env_t *e, uint8_t v) { unsigned c = e->p & FC; e->p = (uint8_t)((e->p & ~FC) | (v & 1)); v = (uint8_t)((v >> 1) | (c << 7)); SETNZ(v); return v; }

static void push(cpu_ctx *c, uint8_t v) { wr(c, (uint16_t)(0x100 | c->e->sp), v); c->e->sp--; }
static uint8_t pull(cpu_ctx *c) { c->e->sp++; return rd(c, (uint16_t)(0x100 | c->e->sp)); }

/* addressing helpers: return effective address, bump *cyc on page cross when `pen` */
static uint16_t ea_zp (cpu_ctx *c, uint16_t pc) { return peek(c, pc); }
static uint16_t ea_zpx(cpu_ctx *c, uint16_t pc) { return (uint8_t)(peek(c, pc) + c->e->x); }
static uint16_t ea_zpy(cpu_ctx *c, uint16_t pc) { return (uint8_t)(peek(c, pc) + c->e->y); }
static uint16_t ea_abs(cpu_ctx *c, uint16_t pc) { return (uint16_t)(peek(c, pc) | (peek(c, (uint16_t)(pc + 1)) << 8)); }
static uint16_t ea_abi(cpu_ctx *c, uint16_t pc, uint8_t idx, int pen, int *cyc)
{
    uint16_t b = ea_abs(c, pc), a = (uint16_t)(b + idx);
    if (pen && ((a ^ b) & 0xFF00)) (*cyc)++;
    return a;
}
static uint16_t ea_inx(cpu_ctx *c, uint16_t pc)
{
    uint8_t z = (uint8_t)(peek(c, pc) + c->e->x);
    return (uint16_t)(peek(c, z) | (peek(c, (uint8_t)(z + 1)) << 8));
}
static uint16_t ea_iny(cpu_ctx *c, uint16_t pc, int pen, int *cyc)
{
    uint8_t z = peek(c, pc);
    uint16_t b = (uint16_t)(peek(c, z) | (peek(c, (uint8_t)(z + 1)) << 8));
    uint16_t a = (uint16_t)(b + c->e->y);
    if (pen && ((a ^ b) & 0xFF00)) (*cyc)++;
    return a;
}

/* addressing mode classes for the table below */
enum { IMP, IMM, ZP, ZPX, ZPY, ABS, ABX, ABY, INX, INY, REL, IND, ACC };
/* access kind: R read, W write, M read-modify-write, N none */
enum { KN, KR, KW, KM };

typedef struct { uint8_t mode, kind, cyc; } opinfo;

/* base cycles (without page-cross / branch penalties) */
static const opinfo OPS[256] = {
/*00*/{IMP,KN,7},{INX,KR,6},{IMP,KN,2},{INX,KM,8},{ZP,KR,3},{ZP,KR,3},{ZP,KM,5},{ZP,KM,5},{IMP,KN,3},{IMM,KR,2},{ACC,KN,2},{IMM,KR,2},{ABS,KR,4},{ABS,KR,4},{ABS,KM,6},{ABS,KM,6},
/*10*/{REL,KN,2},{INY,KR,5},{IMP,KN,2},{INY,KM,8},{ZPX,KR,4},{ZPX,KR,4},{ZPX,KM,6},{ZPX,KM,6},{IMP,KN,2},{ABY,KR,4},{IMP,KN,2},{ABY,KM,7},{ABX,KR,4},{ABX,KR,4},{ABX,KM,7},{ABX,KM,7},
/*20*/{ABS,KN,6},{INX,KR,6},{IMP,KN,2},{INX,KM,8},{ZP,KR,3},{ZP,KR,3},{ZP,KM,5},{ZP,KM,5},{IMP,KN,4},{IMM,KR,2},{ACC,KN,2},{IMM,KR,2},{ABS,KR,4},{ABS,KR,4},{ABS,KM,6},{ABS,KM,6},
/*30*/{REL,KN,2},{INY,KR,5},{IMP,KN,2},{INY,KM,8},{ZPX,KR,4},{ZPX,KR,4},{ZPX,KM,6},{ZPX,KM,6},{IMP,KN,2},{ABY,KR,4},{IMP,KN,2},{ABY,KM,7},{ABX,KR,4},{ABX,KR,4},{ABX,KM,7},{ABX,KM,7},
/*40*/{IMP,KN,6},{INX,KR,6},{IMP,KN,2},{INX,KM,8},{ZP,KR,3},{ZP,KR,3},{ZP,KM,5},{ZP,KM,5},{IMP,KN,3},{IMM,KR,2},{ACC,KN,2},{IMM,KR,2},{ABS,KN,3},{ABS,KR,4},{ABS,KM,6},{ABS,KM,6},
/*50*/{REL,KN,2},{INY,KR,5},{IMP,KN,2},{INY,KM,8},{ZPX,KR,4},{ZPX,KR,4},{ZPX,KM,6},{ZPX,KM,6},{IMP,KN,2},{ABY,KR,4},{IMP,KN,2},{ABY,KM,7},{ABX,KR,4},{ABX,KR,4},{ABX,KM,7},{ABX,KM,7},
/*60*/{IMP,KN,6},{INX,KR,6},{IMP,KN,2},{INX,KM,8},{ZP,KR,3},{ZP,KR,3},{ZP,KM,5},{ZP,KM,5},{IMP,KN,4},{IMM,KR,2},{ACC,KN,2},{IMM,KR,2},{IND,KN,5},{ABS,KR,4},{ABS,KM,6},{ABS,KM,6},
/*70*/{REL,KN,2},{INY,KR,5},{IMP,KN,2},{INY,KM,8},{ZPX,KR,4},{ZPX,KR,4},{ZPX,KM,6},{ZPX,KM,6},{IMP,KN,2},{ABY,KR,4},{IMP,KN,2},{ABY,KM,7},{ABX,KR,4},{ABX,KR,4},{ABX,KM,7},{ABX,KM,7},
/*80*/{IMM,KR,2},{INX,KW,6},{IMM,KR,2},{INX,KW,6},{ZP,KW,3},{ZP,KW,3},{ZP,KW,3},{ZP,KW,3},{IMP,KN,2},{IMM,KR,2},{IMP,KN,2},{IMM,KR,2},{ABS,KW,4},{ABS,KW,4},{ABS,KW,4},{ABS,KW,4},
/*90*/{REL,KN,2},{INY,KW,6},{IMP,KN,2},{INY,KW,6},{ZPX,KW,4},{ZPX,KW,4},{ZPY,KW,4},{ZPY,KW,4},{IMP,KN,2},{ABY,KW,5},{IMP,KN,2},{ABY,KW,5},{ABX,KW,5},{ABX,KW,5},{ABY,KW,5},{ABY,KW,5},
/*A0*/{IMM,KR,2},{INX,KR,6},{IMM,KR,2},{INX,KR,6},{ZP,KR,3},{ZP,KR,3},{ZP,KR,3},{ZP,KR,3},{IMP,KN,2},{IMM,KR,2},{IMP,KN,2},{IMM,KR,2},{ABS,KR,4},{ABS,KR,4},{ABS,KR,4},{ABS,KR,4},
/*B0*/{REL,KN,2},{INY,KR,5},{IMP,KN,2},{INY,KR,5},{ZPX,KR,4},{ZPX,KR,4},{ZPY,KR,4},{ZPY,KR,4},{IMP,KN,2},{ABY,KR,4},{IMP,KN,2},{ABY,KR,4},{ABX,KR,4},{ABX,KR,4},{ABY,KR,4},{ABY,KR,4},
/*C0*/{IMM,KR,2},{INX,KR,6},{IMM,KR,2},{INX,KM,8},{ZP,KR,3},{ZP,KR,3},{ZP,KM,5},{ZP,KM,5},{IMP,KN,2},{IMM,KR,2},{IMP,KN,2},{IMM,KR,2},{ABS,KR,4},{ABS,KR,4},{ABS,KM,6},{ABS,KM,6},
/*D0*/{REL,KN,2},{INY,KR,5},{IMP,KN,2},{INY,KM,8},{ZPX,KR,4},{ZPX,KR,4},{ZPX,KM,6},{ZPX,KM,6},{IMP,KN,2},{ABY,KR,4},{IMP,KN,2},{ABY,KM,7},{ABX,KR,4},{ABX,KR,4},{ABX,KM,7},{ABX,KM,7},
/*E0*/{IMM,KR,2},{INX,KR,6},{IMM,KR,2},{INX,KM,8},{ZP,KR,3},{ZP,KR,3},{ZP,KM,5},{ZP,KM,5},{IMP,KN,2},{IMM,KR,2},{IMP,KN,2},{IMM,KR,2},{ABS,KR,4},{ABS,KR,4},{ABS,KM,6},{ABS,KM,6},
/*F0*/{REL,KN,2},{INY,KR,5},{IMP,KN,2},{INY,KM,8},{ZPX,KR,4},{ZPX,KR,4},{ZPX,KM,6},{ZPX,KM,6},{IMP,KN,2},{ABY,KR,4},{IMP,KN,2},{ABY,KM,7},{ABX,KR,4},{ABX,KR,4},{ABX,KM,7},{ABX,KM,7},
};

static const uint8_t MODE_LEN[13] = { 1, 2, 2, 2, 2, 3, 3, 3, 2, 2, 2, 3, 1 };

/* Execute exactly one instruction under the timing model of DESIGN.md §3:
 *   cycles n are fixed before any data access; every TIA/RIOT/stack/data access of the
 *   instruction happens at cycle n-1; WSYNC stalls after the instruction completes. */
static void exec_one(cpu_ctx *c)
{
    env_t *e = c->e;
    uint16_t pc = e->pc;
    uint8_t op = peek(c, pc);
    opinfo oi = OPS[op];
    int cyc = oi.cyc;
    uint16_t opc = (uint16_t)(pc + 1);
    uint16_t ea = 0;
    int pen = (oi.kind == KR);            /* page-cross penalty applies to reads only */
    uint16_t npc = (uint16_t)(pc + MODE_LEN[oi.mode]);

    switch (oi.mode) {
    case ZP:  ea = ea_zp(c, opc); break;
    case ZPX: ea = ea_zpx(c, opc); break;
    case ZPY: ea = ea_zpy(c, opc); break;
    case ABS: ea = ea_abs(c, opc); break;
    case ABX: ea = ea_abi(c, opc, e->x, pen, &cyc); break;
    case ABY: ea = ea_abi(c, opc, e->y, pen, &cyc); break;
    case INX: ea = ea_inx(c, opc); break;
    case INY: ea = ea_iny(c, opc, pen, &cyc); break;
    case IND: {
        uint16_t p = ea_abs(c, opc);
        uint16_t p2 = (uint16_t)((p & 0xFF00) | ((p + 1) & 0x00FF));   /* page-wrap bug */
        ea = (uint16_t)(peek(c, p) | (peek(c, p2) << 8));
        break; }
    default: break;
    }

    /* branches: resolve penalty before advancing */
    int taken = 0;
    if (oi.mode == REL) {
        uint8_t p = e->p;
        switch (op) {
        case 0x10: taken = !(p & FN); break;
        case 0x30: taken = (p & FN) != 0; break;
        case 0x50: taken = !(p & FV); break;
        case 0x70: taken = (p & FV) != 0; break;
        case 0x90: taken = !(p & FC); break;
        case 0xB0: taken = (p & FC) != 0; break;
        case 0xD0: taken = !(p & FZ); break;
        case 0xF0: taken = (p & FZ) != 0; break;
        }
        if (taken) {
            uint16_t t = (uint16_t)(npc + (int8_t)peek(c, opc));
            cyc += 1 + (((t ^ npc) & 0xFF00) ? 1 : 0);
            npc = t;
        }
    }

    c->wsync = 0;
    advance(c, cyc - 1);

    uint8_t v = 0;
    if (oi.mode == IMM) v = peek(c, opc);
    else if (oi.kind == KR || oi.kind == KM) v = rd(c, ea);

    switch (op) {
    /* ---- loads / stores ---- */
    case 0xA9: case 0xA5: case 0xB5: case 0xAD: case 0xBD: case 0xB9: case 0xA1: case 0xB1: e->a = v; SETNZ(v); break;
    case 0xA2: case 0xA6: case 0xB6: case 0xAE: case 0xBE: e->x = v; SETNZ(v); break;
    case 0xA0: case 0xA4: case 0xB4: case 0xAC: case 0xBC: e->y = v; SETNZ(v); break;
    case 0x85: case 0x95: case 0x8D: case 0x9D: case 0x99: case 0x81: case 0x91: wr(c, ea, e->a); break;
    case 0x86: case 0x96: case 0x8E: wr(c, ea, e->x); break;
    case 0x84: case 0x94: case 0x8C: wr(c, ea, e->y); break;
    /* ---- transfers ---- */
    case 0xAA: e->x = e->a; SETNZ(e->x); break;
    case 0xA8: e->y = e->a; SETNZ(e->y); break;
    case 0x8A: e->a = e->x; SETNZ(e->a); break;
    case 0x98: e->a = e->y; SETNZ(e->a); break;
    case 0xBA: e->x = e->sp; SETNZ(e->x); break;
    case 0x9A: e->sp = e->x; break;
    /* ---- stack ---- */
    case 0x48: push(c, e->a); break;
    case 0x08: push(c, (uint8_t)(e->p | FB | FU)); break;
    case 0x68: e->a = pull(c); SETNZ(e->a); break;
    case 0x28: e->p = (uint8_t)((pull(c) & ~FB) | FU); break;
    /* ---- logic / arithmetic ---- */
    case 0x29: case 0x25: case 0x35: case 0x2D: case 0x3D: case 0x39: case 0x21: case 0x31: e->a &= v; SETNZ(e->a); break;
    case 0x09: case 0x05: case 0x15: case 0x0D: case 0x1D: case 0x19: case 0x01: case 0x11: e->a |= v; SETNZ(e->a); break;
    case 0x49: case 0x45: case 0x55: case 0x4D: case 0x5D: case 0x59: case 0x41: case 0x51: e->a ^= v; SETNZ(e->a); break;
    case 0x69: case 0x65: case 0x75: case 0x6D: case 0x7D: case 0x79: case 0x61: case 0x71: op_adc(e, v); break;
    case 0xE9: case 0xEB: case 0xE5: case 0xF5: case 0xED: case 0xFD: case 0xF9: case 0xE1: case 0xF1: op_sbc(e, v); break;
    case 0xC9: case 0xC5: case 0xD5: case 0xCD: case 0xDD: case 0xD9: case 0xC1: case 0xD1: op_cmp(e, e->a, v); break;
    case 0xE0: case 0xE4: case 0xEC: op_cmp(e, e->x, v); break;
    case 0xC0: case 0xC4: case 0xCC: op_cmp(e, e->y, v); break;
    case 0x24: case 0x2C:
        e->p = (uint8_t)((e->p & ~(FN | FV | FZ)) | (v & 0xC0) | ((v & e->a) ? 0 : FZ)); break;
    /* ---- shifts ---- */
    case 0x0A: e->a = op_asl(e, e->a); break;
    case 0x4A: e->a = op_lsr(e, e->a); break;
    case 0x2A: e->a = op_rol(e, e->a); break;
    case 0x6A: e->a = op_ror(e, e->a); break;
    case 0x06: case 0x16: case 0x0E: case 0x1E: wr(c, ea, op_asl(e, v)); break;
    case 0x46: case 0x56: case 0x4E: case 0x5E: wr(c, ea, op_lsr(e, v)); break;
    case 0x26: case 0x36: case 0x2E: case 0x3E: wr(c, ea, op_rol(e, v)); break;
    case 0x66: case 0x76: case 0x6E: case 0x7E: wr(c, ea, op_ror(e, v)); break;
    /* ---- inc / dec ---- */
    case 0xE6: case 0xF6: case 0xEE: case 0xFE: v++; SETNZ(v); wr(c, ea, v); break;
    case 0xC6: case 0xD6: case 0xCE: case 0xDE: v--; SETNZ(v); wr(c, ea, v); break;
    case 0xE8: e->x++; SETNZ(e->x); break;
    case 0xC8: e->y++; SETNZ(e->y); break;
    case 0xCA: e->x--; SETNZ(e->x); break;
    case 0x88: e->y--; SETNZ(e->y); break;
    /* ---- flags ---- */
    case 0x18: e->p &= (uint8_t)~FC; break;
    case 0x38: e->p |= FC; break;
    case 0x58: e->p &= (uint8_t)~FI; break;
    case 0x78: e->p |= FI; break;
    case 0xB8: e->p &= (uint8_t)~FV; break;
    case 0xD8: e->p &= (uint8_t)~FD; break;
    case 0xF8: e->p |= FD; break;
    /* ---- control flow ---- */
    case 0x4C: npc = ea; break;
    case 0x6C: npc = ea; break;
    case 0x20: {
        uint16_t ret = (uint16_t)(pc + 2);
        push(c, (uint8_t)(ret >> 8)); push(c, (uint8_t)ret);
        npc = ea; break; }
    case 0x60: { uint8_t lo = pull(c); uint8_t hi = pull(c); npc = (uint16_t)(((hi << 8) | lo) + 1); break; }
    case 0x40: {
        e->p = (uint8_t)((pull(c) & ~FB) | FU);
        uint8_t lo = pull(c); uint8_t hi = pull(c); npc = (uint16_t)((hi << 8) | lo); break; }
    case 0x00: {
        uint16_t ret = (uint16_t)(pc + 2);
        push(c, (uint8_t)(ret >> 8)); push(c, (uint8_t)ret);
        push(c, (uint8_t)(e->p | FB | FU));
        e->p |= FI;
        npc = (uint16_t)(peek(c, 0xFFFE) | (peek(c, 0xFFFF) << 8)); break; }
    /* ---- stable undocumented ---- */
    case 0xA7: case 0xB7: case 0xAF: case 0xBF: case 0xA3: case 0xB3: e->a = e->x = v; SETNZ(v); break;     /* LAX */
    case 0x87: case 0x97: case 0x8F: case 0x83: wr(c, ea, (uint8_t)(e->a & e->x)); break;                   /* SAX */
    case 0xC7: case 0xD7: case 0xCF: case 0xDF: case 0xDB: case 0xC3: case 0xD3:                             /* DCP */
        v--; wr(c, ea, v); op_cmp(e, e->a, v); break;
    case 0xE7: case 0xF7: case 0xEF: case 0xFF: case 0xFB: case 0xE3: case 0xF3:                             /* ISC */
        v++; wr(c, ea, v); op_sbc(e, v); break;
    case 0x07: case 0x17: case 0x0F: case 0x1F: case 0x1B: case 0x03: case 0x13:                             /* SLO */
        v = op_asl(e, v); wr(c, ea, v); e->a |= v; SETNZ(e->a); break;
    case 0x27: case 0x37: case 0x2F: case 0x3F: case 0x3B: case 0x23: case 0x33:                             /* RLA */
        v = op_rol(e, v); wr(c, ea, v); e->a &= v; SETNZ(e->a); break;
    case 0x47: case 0x57: case 0x4F: case 0x5F: case 0x5B: case 0x43: case 0x53:                             /* SRE */
        v = op_lsr(e, v); wr(c, ea, v); e->a ^= v; SETNZ(e->a); break;
    case 0x67: case 0x77: case 0x6F: case 0x7F: case 0x7B: case 0x63: case 0x73:                             /* RRA */
        v = op_ror(e, v); wr(c, ea, v); op_adc(e, v); break;
    case 0x0B: case 0x2B: e->a &= v; SETNZ(e->a); e->p = (uint8_t)((e->p & ~FC) | (e->a >> 7)); break;      /* ANC */
    case 0x4B: e->a &= v; e->a = op_lsr(e, e->a); break;                                                     /* ALR */
    case 0xCB: {                                                                                             /* SBX */
        unsigned t = (unsigned)(e->a & e->x) - v;
        e->p &= (uint8_t)~FC; if (!(t & 0x100)) e->p |= FC;
        e->x = (uint8_t)t; SETNZ(e->x); break; }
    default:
        /* documented NOP (0xEA), undocumented NOPs, and every unstable/JAM opcode: no effect
           beyond the (already performed) operand read and the table's length / cycles */
        break;
    }

    e->pc = npc;
    advance(c, 1);
    if (c->wsync && e->hcyc != 0) advance(c, STM_CYC_PER_LINE - e->hcyc);
}

/* ------------------------------------------------------------------------------------------- */
/* frame / agent-step level                                                                     */
/* ------------------------------------------------------------------------------------------- */
void orc_env_reset(const orc_game *g, uint8_t *blob)
{
    env_t *e = (env_t *)blob;
    memset(e, 0, sizeof(*e));
    e->sp = 0xFD; e->p = FI | FU;
    e->tshift = 10;
    e->paddle[0] = e->paddle[1] = STM_PADDLE_INIT;
    cpu_ctx c = { g, e, NULL, 0, 0 };
    e->pc = (uint16_t)(peek(&c, 0xFFFC) | (peek(&c, 0xFFFD) << 8));
}

void orc_exec_one(const orc_game *g, uint8_t *blob)
{
    cpu_ctx c = { g, (env_t *)blob, NULL, 0, 0 };
    exec_one(&c);
}

static void set_action(const orc_game *g, env_t *e, int ale_action)
{
    e->action = (uint8_t)ale_action;
    if (g->d.controller == STM_CTRL_PADDLE) {
        int up, down, left, right;
        action_dirs(e->action, &up, &down, &left, &right);
        int p = e->paddle[0];
        if (right) p -= STM_PADDLE_DELTA;
        if (left)  p += STM_PADDLE_DELTA;
        if (p < STM_PADDLE_MIN) p = STM_PADDLE_MIN;
        if (p > STM_PADDLE_MAX) p = STM_PADDLE_MAX;
        e->paddle[0] = (uint8_t)p;
    }
    update_latch(e, g);
}

/* one emulated frame; fb (160x210) may be NULL => collisions only */
void orc_frame(const orc_game *g, uint8_t *blob, int ale_action, uint8_t *fb)
{
    env_t *e = (env_t *)blob;
    cpu_ctx c = { g, e, fb, 0, 0 };
    set_action(g, e, ale_action);
    e->flags &= (uint8_t)~STM_F_FRAME_DONE;
    while (!(e->flags & STM_F_FRAME_DONE)) exec_one(&c);
    /* rows the frame never reached (VSYNC came at scanline L < YSTART+210) are black, and so is
       the partially drawn row L-YSTART */
    if (fb) {
        int r0 = c.end_line_no - STM_YSTART; if (r0 < 0) r0 = 0;
        if (r0 < STM_VIS_H) memset(fb + r0 * STM_VIS_W, 0, (size_t)(STM_VIS_H - r0) * STM_VIS_W);
    }
    e->frame++;
}

static int bcd_score(const env_t *e, const uint8_t addr[3])
{
    int s = 0;
    for (int i = 0; i < 3; i++) {
        if (addr[i] == 0xFF) continue;
        uint8_t b = e->ram[addr[i] & 0x7F];
        s = s * 100 + (b >> 4) * 10 + (b & 15);
    }
    return s;
}

int orc_lives(const orc_game *g, const uint8_t *blob)
{
    const env_t *e = (const env_t *)blob;
    if (g->d.lives_addr == 0xFF) return 0;
    return (e->ram[g->d.lives_addr & 0x7F] >> g->d.lives_shift) & g->d.lives_mask;
}

int orc_terminal(const orc_game *g, const uint8_t *blob)
{
    const env_t *e = (const env_t *)blob;
    if (g->d.term_addr == 0xFF) return 0;
    return (e->ram[g->d.term_addr & 0x7F] & g->d.term_mask) == g->d.term_value;
}

/* restates torchcule get_data(episodic_life, weight, done, rewards, lives) as used at
   cule_bfs.py:157-158: rewards += weight*r_frame; done |= terminal | life lost; lives updated */
void orc_get_data(const orc_game *g, uint8_t *blob, int episodic_life, float weight,
                  uint8_t *done, float *reward, int32_t *lives)
{
    env_t *e = (env_t *)blob;
    int score = bcd_score(e, g->d.score_addr) - bcd_score(e, g->d.score2_addr);
    int r = score - e->score_prev;
    e->score_prev = score;
    volatile float wr_ = weight * (float)r;      /* separate multiply and add: no FMA contraction */
    *reward = *reward + wr_;
    int lv = orc_lives(g, blob);
    int term = 0;
    if (g->d.term_addr != 0xFF)
        term = (e->ram[g->d.term_addr & 0x7F] & g->d.term_mask) == g->d.term_value;
    if (term || (episodic_life && lv < *lives)) *done = 1;
    *lives = lv;
}

/* grey + bilinear 160x210 -> 84x84, integer arithmetic */
void orc_render_obs(const uint8_t *fb, uint8_t *out)
{
    for (int oy = 0; oy < STM_OBS_H; oy++) {
        int ny = 10 * oy + 3, y0 = ny >> 2, fy = ny & 3;
        int y1 = y0 + 1 < STM_VIS_H ? y0 + 1 : STM_VIS_H - 1;
        for (int ox = 0; ox < STM_OBS_W; ox++) {
            int nx = 80 * ox + 19, x0 = nx / 42, fx = nx % 42;
            int x1 = x0 + 1 < STM_VIS_W ? x0 + 1 : STM_VIS_W - 1;
            int p00 = GRAY[fb[y0 * STM_VIS_W + x0] & 127], p01 = GRAY[fb[y0 * STM_VIS_W + x1] & 127];
            int p10 = GRAY[fb[y1 * STM_VIS_W + x0] & 127], p11 = GRAY[fb[y1 * STM_VIS_W + x1] & 127];
            int top = (42 - fx) * p00 + fx * p01, bot = (42 - fx) * p10 + fx * p11;
            out[oy * STM_OBS_W + ox] = (uint8_t)(((4 - fy) * top + fy * bot + 84) / 168);
        }
    }
}

/* One agent step = `frameskip` emulated frames with get_data after each (cule_bfs.py:152-158),
 * observation = max of the grey 84x84 renders of the last two frames (:162-168).
 * obs may be NULL.  scratch must hold 2*STM_NATIVE_BYTES. A done env is frozen. */
void orc_agent_step(const orc_game *g, uint8_t *blob, int policy_action, int frameskip,
                    int episodic_life, float weight, uint8_t *done, float *reward, int32_t *lives,
                    uint8_t *obs, uint8_t *scratch)
{
    uint8_t *fa = scratch, *fbuf = scratch + STM_NATIVE_BYTES;
    int ale = g->d.actions[policy_action];
    if (*done) {
        /* frozen env (done is tested once, at the start of the agent step): no emulation; its new
           observation is all-zero (the reference zeroes observations1/2 up front,
           cule_bfs.py:116-117, and a skipped env writes nothing) */
        if (obs) memset(obs, 0, STM_OBS_BYTES);
        return;
    }
    for (int f = 0; f < frameskip; f++) {
        uint8_t *target = NULL;
        if (obs && f == frameskip - 2) target = fa;
        if (obs && f == frameskip - 1) target = fbuf;
        orc_frame(g, blob, ale, target);
        orc_get_data(g, blob, episodic_life, weight, done, reward, lives);
    }
    if (obs) {
        uint8_t o1[STM_OBS_BYTES], o2[STM_OBS_BYTES];
        if (frameskip >= 2) orc_render_obs(fa, o2); else memset(o2, 0, sizeof o2);
        orc_render_obs(fbuf, o1);
        for (int i = 0; i < STM_OBS_BYTES; i++) obs[i] = o1[i] > o2[i] ? o1[i] : o2[i];
    }
}

/*
 * Breadth-first expansion of R roots to depth d (cule_bfs.py:98-209, batched over roots).
 *   roots      R x 256 state blobs                       (cule_bfs.py:110-112)
 *   root_stack R x 4 x 84 x 84 u8                        (cule_bfs.py:99)
 * outputs (L = A^d):
 *   leaf_stack R x L x 4 x 84 x 84 u8 ; leaf_reward R x L f32 (sign() applied if clip_reward,
 *   :205-206) ; leaf_done R x L u8 ; leaf_lives R x L i32 ; leaf_state R x L x 256 (may be NULL)
 * Child i of a level has parent i / A and takes action i % A (:90,144).
 * Returns 0, or -1 on allocation failure.
 */
int orc_bfs(const orc_game *g, const uint8_t *roots, const uint8_t *root_stack, int R, int depth,
            float gamma, int frameskip, int episodic_life, int clip_reward,
            uint8_t *leaf_stack, float *leaf_reward, uint8_t *leaf_done, int32_t *leaf_lives,
            uint8_t *leaf_state, int nthreads)
{
    const int A = g->d.n_actions;
    size_t n = (size_t)R, nmax = (size_t)R;
    for (int k = 0; k < depth; k++) nmax *= (size_t)A;
    const size_t SB = STM_STACK * STM_OBS_BYTES;
    uint8_t *st[2], *ob[2], *dn[2]; float *rw[2]; int32_t *lv[2];
    for (int i = 0; i < 2; i++) {
        st[i] = malloc(nmax * STM_STATE_BYTES); ob[i] = malloc(nmax * SB); dn[i] = malloc(nmax);
        rw[i] = malloc(nmax * sizeof(float)); lv[i] = malloc(nmax * sizeof(int32_t));
        if (!st[i] || !ob[i] || !dn[i] || !rw[i] || !lv[i]) return -1;
    }
    memcpy(st[0], roots, n * STM_STATE_BYTES);
    memcpy(ob[0], root_stack, n * SB);
    for (size_t i = 0; i < n; i++) { rw[0][i] = 0.f; dn[0][i] = 0; lv[0][i] = orc_lives(g, roots + i * STM_STATE_BYTES); }
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    int cur = 0;
    for (int k = 0; k < depth; k++) {
        size_t nn = n * (size_t)A;
        int nxt = cur ^ 1;
        float w = (float)pow((double)gamma, (double)k);     /* gamma ** depth (Python double) narrowed */
        #pragma omp parallel
        {
            uint8_t *scratch = malloc(2 * STM_NATIVE_BYTES);
            uint8_t obs[STM_OBS_BYTES];
            #pragma omp for schedule(dynamic, 8)
            for (long long ii = 0; ii < (long long)nn; ii++) {
                size_t i = (size_t)ii, p = i / (size_t)A; int a = (int)(i % (size_t)A);
                memcpy(st[nxt] + i * STM_STATE_BYTES, st[cur] + p * STM_STATE_BYTES, STM_STATE_BYTES);
                rw[nxt][i] = rw[cur][p]; dn[nxt][i] = dn[cur][p]; lv[nxt][i] = lv[cur][p];
                orc_agent_step(g, st[nxt] + i * STM_STATE_BYTES, a, frameskip, episodic_life, w,
                               &dn[nxt][i], &rw[nxt][i], &lv[nxt][i], obs, scratch);
                memcpy(ob[nxt] + i * SB, ob[cur] + p * SB + STM_OBS_BYTES, SB - STM_OBS_BYTES);
                memcpy(ob[nxt] + i * SB + SB - STM_OBS_BYTES, obs, STM_OBS_BYTES);
            }
            free(scratch);
        }
        n = nn; cur = nxt;
    }
    memcpy(leaf_stack, ob[cur], n * SB);
    for (size_t i = 0; i < n; i++) {
        float r = rw[cur][i];
        if (clip_reward) r = (r > 0.f) - (r < 0.f);
        leaf_reward[i] = r; leaf_done[i] = dn[cur][i]; leaf_lives[i] = lv[cur][i];
    }
    if (leaf_state) memcpy(leaf_state, st[cur], n * STM_STATE_BYTES);
    for (int i = 0; i < 2; i++) { free(st[i]); free(ob[i]); free(dn[i]); free(rw[i]); free(lv[i]); }
    return 0;
}
