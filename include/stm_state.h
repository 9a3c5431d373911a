/*
 * stm_state.h — on-device / on-host layout of one emulator environment ("env state blob")
 * and of a game descriptor.  This is a DATA FORMAT, shared by the CUDA path and by the CPU
 * oracle so that blobs can be compared byte for byte.
 *
 * Replaces (reference side): the torchcule env tensors `states`, `ram`, `frame_states`
 * that policies/cule_bfs.py:73-77,83-96,110-112 copies around.  In the reference those are three
 * tensors of CuLE-private layout; here they are one 256-byte row per env:
 *     ram          = blob[  0:128]   (RIOT RAM)
 *     states       = blob[128:256]   (6507 + RIOT + TIA registers, timing, input, score latch)
 *     frame_states = (nothing: the TIA needs no separate render state; zero-width view)
 * `rewards` (f32), `done` (u8) and `lives` (i32) stay separate arrays as in the reference
 * (cule_bfs.py:85-87,94-96).
 *
 * All multi-byte fields are little-endian and naturally aligned.
 */
#ifndef STM_STATE_H
#define STM_STATE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STM_STATE_BYTES   256   /* S in DESIGN.md / SURVEY §8(d) */
#define STM_RAM_BYTES     128

/* ---- timing model ------------------------------------------------------------------------ */
#define STM_CYC_PER_LINE  76    /* CPU cycles per scanline (228 colour clocks / 3)            */
#define STM_HBLANK_CLK    68    /* colour clocks of horizontal blank                          */
#define STM_VIS_W         160   /* visible pixels per line                                    */
#define STM_VIS_H         210   /* rows of the native frame handed to the frame kernel        */
#define STM_YSTART        34    /* first scanline (counted from VSYNC-on) that is row 0       */
#define STM_MAX_LINES     320   /* a frame is force-ended here if the ROM never strobes VSYNC */
#define STM_NATIVE_BYTES  (STM_VIS_W * STM_VIS_H)   /* 33 600 B, one 7-bit colour index/pixel */
#define STM_OBS_W         84
#define STM_OBS_H         84
#define STM_OBS_BYTES     (STM_OBS_W * STM_OBS_H)   /* 7 056 */
#define STM_STACK         4

/* ---- paddle model (ALE-like: LEFT/RIGHT move the pot a fixed amount per emulated frame) ---- */
#define STM_PADDLE_MIN    20    /* scanlines after dump release until INPTx bit7 goes high    */
#define STM_PADDLE_MAX    190
#define STM_PADDLE_DELTA  5
#define STM_PADDLE_INIT   105

/* ---- flags byte --------------------------------------------------------------------------- */
#define STM_F_HMOVE_BLANK 0x01  /* HMOVE strobed during HBLANK of the current line (8-px comb) */
#define STM_F_FRAME_DONE  0x02  /* set by VSYNC 0->1 or by the STM_MAX_LINES cap                */

/* object order used by pos[] */
enum { STM_P0 = 0, STM_P1 = 1, STM_M0 = 2, STM_M1 = 3, STM_BL = 4 };

typedef struct stm_env_state {
    uint8_t  ram[STM_RAM_BYTES];   /*   0 */
    /* 6507 */
    uint16_t pc;                   /* 128 */
    uint8_t  a, x, y, sp, p;       /* 130..134 */
    uint8_t  bank;                 /* 135  F8: 0/1 */
    /* timing */
    uint16_t line;                 /* 136  scanline since VSYNC-on */
    uint8_t  hcyc;                 /* 138  CPU cycle inside the line, 0..75 */
    uint8_t  rx;                   /* 139  pixels [0,rx) of this line are already rendered */
    uint32_t frame;                /* 140  emulated frames since reset */
    /* RIOT 6532 */
    int32_t  timer;                /* 144  cycles until the interval timer reaches zero (can go <0) */
    uint8_t  tshift;               /* 148  0,3,6,10 */
    uint8_t  tflag;                /* 149  bit7: timer underflow latch (INSTAT) */
    uint8_t  swcha_out, swacnt;    /* 150,151 */
    uint8_t  swchb_out, swbcnt;    /* 152,153 */
    /* input */
    uint8_t  action;               /* 154  ALE action code held for this frame */
    uint8_t  flags;                /* 155  STM_F_* */
    uint16_t cx;                   /* 156  collision latches: bit 2r = D6, bit 2r+1 = D7 of CX reg r */
    uint8_t  charge;               /* 158  scanlines since paddle dump released (saturates 255) */
    uint8_t  inpt_latch;           /* 159  bit0: INPT4 latched low */
    uint8_t  paddle[2];            /* 160  pot positions, STM_PADDLE_MIN..MAX */
    uint8_t  pos[5];               /* 162  first-copy start pixel of P0,P1,M0,M1,BL (0..159) */
    uint8_t  pad0;                 /* 167 */
    /* TIA write-side registers */
    uint8_t  vsync, vblank;        /* 168 */
    uint8_t  nusiz0, nusiz1;       /* 170 */
    uint8_t  colup0, colup1, colupf, colubk; /* 172 */
    uint8_t  ctrlpf, refp0, refp1; /* 176 */
    uint8_t  pf0, pf1, pf2;        /* 179 */
    uint8_t  grp0, grp1, grp0_old, grp1_old; /* 182 */
    uint8_t  enam0, enam1, enabl, enabl_old; /* 186 */
    uint8_t  hmp0, hmp1, hmm0, hmm1, hmbl;   /* 190 */
    uint8_t  vdelp0, vdelp1, vdelbl;         /* 195 */
    uint8_t  resmp0, resmp1;       /* 198 */
    /* game latch */
    int32_t  score_prev;           /* 200  score decoded after the previous frame */
    uint8_t  reserved[52];         /* 204..255, always zero */
} stm_env_state;

/* ---- game descriptor (what torchcule.atari.Rom + CuLE's per-game settings supply,
 *      policies/cule_bfs.py:29-31) ------------------------------------------------------------ */
enum { STM_BANK_NONE = 0, STM_BANK_F8 = 1 };
enum { STM_CTRL_JOYSTICK = 0, STM_CTRL_PADDLE = 1 };

typedef struct stm_game_desc {
    uint32_t rom_size;        /* 2048, 4096 (STM_BANK_NONE) or 8192 (STM_BANK_F8) */
    uint8_t  bank_scheme;
    uint8_t  controller;
    uint8_t  n_actions;       /* size of the minimal action set */
    uint8_t  actions[18];     /* ALE action codes, index = policy action */
    uint8_t  score_addr[3];   /* RAM index (0..127) of BCD bytes, most significant first; 0xFF unused */
    uint8_t  score2_addr[3];  /* opponent score (subtracted); 0xFF unused */
    uint8_t  lives_addr;      /* RAM index; 0xFF => lives always 0 */
    uint8_t  lives_shift, lives_mask;
    uint8_t  term_addr;       /* terminal <=> (ram[term_addr] & term_mask) == term_value; 0xFF => never */
    uint8_t  term_mask, term_value;
    uint8_t  pad[3];
} stm_game_desc;

#ifdef __cplusplus
}
#endif
#endif /* STM_STATE_H */
