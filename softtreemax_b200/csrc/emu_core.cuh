// emu_core.cuh — Atari 2600 (6507 + RIOT + TIA) core used by the batched step kernel (K2).
//
// One instance of Emu<M> is one environment, owned by one GPU thread.  The env state blob
// (include/stm_state.h) lives in shared memory, word-interleaved across the block so that every
// lane owns one bank; `M` is the memory policy that hides that layout.  The same source compiles
// with a plain-array policy on the host, which is how tests/ cross-check it against the oracle
// without a GPU (tests/tools/host_core.cpp) — the host build is a test aid, never a product path.
//
// Structure (deliberately unlike the oracle's big switch):
//   - opcodes are decoded through a 256-entry table into (addressing mode, access kind, base
//     cycles, operation class); lanes that run different opcodes with the same addressing mode or
//     the same operation class share code, which bounds SIMT divergence to ~13 + ~60 paths;
//   - the TIA is evaluated a 32-pixel word at a time from per-object bit masks (collisions are
//     word ANDs, colours are layered byte selects), not pixel by pixel;
//   - instruction fetch reads 4 ROM bytes with two aligned word loads and a funnel shift.
//
// Timing model: DESIGN.md §3 (identical to oracle/stm_oracle.c, which is the specification).
// Reference call sites this replaces: policies/cule_bfs.py:154-158 (torchcule step + get_data).
#pragma once
#include <stdint.h>
// hides a value's origin from the optimiser (no code is emitted)
#define STM_OPAQUE(v) asm volatile("" : "+r"(v))
// branch-weight hints: they steer block placement so that the common path of one emulated instruction
// is fall-through code (the step kernel is instruction-fetch sensitive: a 6 KB L0 I-cache per scheduler)
#define STM_LIKELY(c) __builtin_expect(!!(c), 1)
#define STM_UNLIKELY(c) __builtin_expect(!!(c), 0)
#include "../../include/stm_state.h"

#ifdef __CUDACC__
#define STM_DEV __device__ __forceinline__          /* member functions */
#define STM_FN  __device__ __forceinline__          /* free functions   */
#define STM_NOINLINE static __device__ __noinline__
#define STM_WARP_ALL(mask, pred) __all_sync((mask), (pred))
#define STM_WARP_MIN(mask, v) __reduce_min_sync((mask), (v))
// Scanline lock-step scope: the warps of a thread block that share a scheduler (warp id mod 4) and with it an L0 instruction
// cache — one named barrier per group, `group` = 32 x the group's warps | barrier id << 16.  Walking the same code region together is what
// keeps the group's instruction fetches in L0 (K2 is bound by instruction supply: DESIGN.md §5).  Every warp of a group must keep
// calling this until it returns true; stm_step.cuh keeps dead lanes in the loop for that reason.
static __device__ __forceinline__ bool stm_group_all(uint32_t group, bool pred)     // group = threads | barrier id << 16
{
    const uint32_t w = __all_sync(0xFFFFFFFFu, pred);
    if ((group & 0xFFFFu) == 32u) return w != 0;
    uint32_t r;
    asm volatile("{\n .reg .pred p, q;\n setp.ne.u32 q, %1, 0;\n barrier.red.and.pred p, %2, %3, q;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(r) : "r"(w), "r"(group >> 16), "r"(group & 0xFFFFu) : "memory");
    return r != 0;
}
#define STM_LINE_ALL(group, pred) stm_group_all((group), (pred))
#else
#define STM_LINE_ALL(mask, pred) (pred)
#define STM_DEV inline
#define STM_FN  static inline
#define STM_NOINLINE static
#define STM_WARP_ALL(mask, pred) (pred)
#define STM_WARP_MIN(mask, v) (v)
#endif

namespace stm {

// ---- decode table ----------------------------------------------------------------------------
enum Mode : uint32_t { M_IMP, M_IMM, M_ZP, M_ZPX, M_ZPY, M_ABS, M_ABX, M_ABY, M_INX, M_INY, M_REL, M_IND, M_ACC };
enum Kind : uint32_t { K_N, K_R, K_W, K_M };
enum Cls : uint32_t {
    C_NOP, C_LDA, C_LDX, C_LDY, C_LAX, C_STA, C_STX, C_STY, C_SAX, C_TAX, C_TAY, C_TXA, C_TYA, C_TSX,
    C_TXS, C_PHA, C_PHP, C_PLA, C_PLP, C_AND, C_ORA, C_EOR, C_ADC, C_SBC, C_CMP, C_CPX, C_CPY, C_BIT,
    C_ASL, C_LSR, C_ROL, C_ROR, C_INC, C_DEC, C_INX, C_INY, C_DEX, C_DEY, C_CLC, C_SEC, C_CLI, C_SEI,
    C_CLV, C_CLD, C_SED, C_JMP, C_JSR, C_RTS, C_RTI, C_BRK, C_BRA, C_DCP, C_ISC, C_SLO, C_RLA, C_SRE,
    C_RRA, C_ANC, C_ALR, C_SBX, C_COUNT
};
// entry = mode | kind<<4 | cycles<<6 | cls<<10
static inline uint32_t pack_op(uint32_t mode, uint32_t kind, uint32_t cyc, uint32_t cls)
{ return mode | (kind << 4) | (cyc << 6) | (cls << 10); }

// Builds the 256-entry decode table from the regular structure of the 6502 opcode matrix
// (op = aaabbbcc) plus the irregular column cc=00 and the single-byte instructions.
static inline void build_decode_table(uint32_t *t)
{
    auto cyc_of = [](uint32_t mode, uint32_t kind) -> uint32_t {
        switch (mode) {
        case M_IMM: return 2;
        case M_ZP:  return kind == K_M ? 5 : 3;
        case M_ZPX: case M_ZPY: return kind == K_M ? 6 : 4;
        case M_ABS: return kind == K_M ? 6 : 4;
        case M_ABX: case M_ABY: return kind == K_M ? 7 : (kind == K_W ? 5 : 4);
        case M_INX: return kind == K_M ? 8 : 6;
        case M_INY: return kind == K_M ? 8 : (kind == K_W ? 6 : 5);
        default: return 2;
        }
    };
    auto set = [&](uint32_t op, uint32_t cls, uint32_t mode, uint32_t kind, int cyc = -1) {
        t[op] = pack_op(mode, kind, cyc < 0 ? cyc_of(mode, kind) : (uint32_t)cyc, cls);
    };
    static const uint32_t MODE01[8] = { M_INX, M_ZP, M_IMM, M_ABS, M_INY, M_ZPX, M_ABY, M_ABX };
    for (uint32_t op = 0; op < 256; op++) set(op, C_NOP, M_IMP, K_N, 2);
    for (uint32_t aaa = 0; aaa < 8; aaa++) for (uint32_t bbb = 0; bbb < 8; bbb++) {
        // cc = 01: accumulator ALU group
        {
            static const uint32_t CL[8] = { C_ORA, C_AND, C_EOR, C_ADC, C_STA, C_LDA, C_CMP, C_SBC };
            uint32_t op = (aaa << 5) | (bbb << 2) | 1, cls = CL[aaa], mode = MODE01[bbb];
            if (cls == C_STA && mode == M_IMM) set(op, C_NOP, M_IMM, K_R);
            else set(op, cls, mode, cls == C_STA ? K_W : K_R);
        }
        // cc = 11: undocumented combinations of the two groups
        {
            static const uint32_t CL[8] = { C_SLO, C_RLA, C_SRE, C_RRA, C_SAX, C_LAX, C_DCP, C_ISC };
            uint32_t op = (aaa << 5) | (bbb << 2) | 3, cls = CL[aaa], mode = MODE01[bbb];
            if (mode == M_IMM) {
                static const uint32_t IMMCL[8] = { C_ANC, C_ANC, C_ALR, C_NOP, C_NOP, C_NOP, C_SBX, C_SBC };
                set(op, IMMCL[aaa], M_IMM, K_R);
            } else if (cls == C_SAX) {
                if (mode == M_ZPX) mode = M_ZPY;
                if (mode == M_ABX) mode = M_ABY;
                bool real = (bbb == 0 || bbb == 1 || bbb == 3 || bbb == 5);
                set(op, real ? C_SAX : C_NOP, mode, K_W);       // 0x93/0x9B/0x9F: unstable -> no effect
            } else if (cls == C_LAX) {
                if (mode == M_ZPX) mode = M_ZPY;
                if (mode == M_ABX) mode = M_ABY;
                set(op, bbb == 6 ? C_NOP : C_LAX, mode, K_R);   // 0xBB LAS -> read only
            } else set(op, cls, mode, K_M);
        }
        // cc = 10: read-modify-write group and the X register loads/stores
        {
            static const uint32_t CL[8] = { C_ASL, C_ROL, C_LSR, C_ROR, C_STX, C_LDX, C_DEC, C_INC };
            uint32_t op = (aaa << 5) | (bbb << 2) | 2, cls = CL[aaa];
            bool xreg = (cls == C_STX || cls == C_LDX);
            switch (bbb) {
            case 0: if (cls == C_LDX) set(op, C_LDX, M_IMM, K_R); else if (aaa >= 4) set(op, C_NOP, M_IMM, K_R); break;
            case 1: set(op, cls, M_ZP, cls == C_STX ? K_W : (cls == C_LDX ? K_R : K_M)); break;
            case 2: if (aaa < 4) set(op, cls, M_ACC, K_N, 2); break;
            case 3: set(op, cls, M_ABS, cls == C_STX ? K_W : (cls == C_LDX ? K_R : K_M)); break;
            case 5: set(op, cls, xreg ? M_ZPY : M_ZPX, cls == C_STX ? K_W : (cls == C_LDX ? K_R : K_M)); break;
            case 7:
                if (cls == C_STX) set(op, C_NOP, M_ABY, K_W);    // 0x9E SHX: no effect
                else set(op, cls, xreg ? M_ABY : M_ABX, cls == C_LDX ? K_R : K_M);
                break;
            default: break;                                       // bbb 4: JAM, bbb 6: NOP/TXS/TSX below
            }
        }
    }
    // column cc = 00
    static const uint32_t ZPNOP[] = { 0x04, 0x44, 0x64 }, ZPXNOP[] = { 0x14, 0x34, 0x54, 0x74, 0xD4, 0xF4 },
                          ABXNOP[] = { 0x1C, 0x3C, 0x5C, 0x7C, 0xDC, 0xFC };
    for (uint32_t o : ZPNOP) set(o, C_NOP, M_ZP, K_R);
    for (uint32_t o : ZPXNOP) set(o, C_NOP, M_ZPX, K_R);
    for (uint32_t o : ABXNOP) set(o, C_NOP, M_ABX, K_R);
    set(0x0C, C_NOP, M_ABS, K_R); set(0x80, C_NOP, M_IMM, K_R);
    set(0x9C, C_NOP, M_ABX, K_W);                                 // SHY: no effect
    set(0x24, C_BIT, M_ZP, K_R);  set(0x2C, C_BIT, M_ABS, K_R);
    set(0x84, C_STY, M_ZP, K_W);  set(0x8C, C_STY, M_ABS, K_W); set(0x94, C_STY, M_ZPX, K_W);
    set(0xA0, C_LDY, M_IMM, K_R); set(0xA4, C_LDY, M_ZP, K_R);  set(0xAC, C_LDY, M_ABS, K_R);
    set(0xB4, C_LDY, M_ZPX, K_R); set(0xBC, C_LDY, M_ABX, K_R);
    set(0xC0, C_CPY, M_IMM, K_R); set(0xC4, C_CPY, M_ZP, K_R);  set(0xCC, C_CPY, M_ABS, K_R);
    set(0xE0, C_CPX, M_IMM, K_R); set(0xE4, C_CPX, M_ZP, K_R);  set(0xEC, C_CPX, M_ABS, K_R);
    set(0x4C, C_JMP, M_ABS, K_N, 3); set(0x6C, C_JMP, M_IND, K_N, 5);
    set(0x20, C_JSR, M_ABS, K_N, 6); set(0x60, C_RTS, M_IMP, K_N, 6);
    set(0x40, C_RTI, M_IMP, K_N, 6); set(0x00, C_BRK, M_IMP, K_N, 7);
    for (uint32_t i = 0; i < 8; i++) set(0x10 | (i << 5), C_BRA, M_REL, K_N, 2);
    set(0x08, C_PHP, M_IMP, K_N, 3); set(0x28, C_PLP, M_IMP, K_N, 4);
    set(0x48, C_PHA, M_IMP, K_N, 3); set(0x68, C_PLA, M_IMP, K_N, 4);
    set(0x88, C_DEY, M_IMP, K_N); set(0xA8, C_TAY, M_IMP, K_N); set(0xC8, C_INY, M_IMP, K_N); set(0xE8, C_INX, M_IMP, K_N);
    set(0x18, C_CLC, M_IMP, K_N); set(0x38, C_SEC, M_IMP, K_N); set(0x58, C_CLI, M_IMP, K_N); set(0x78, C_SEI, M_IMP, K_N);
    set(0x98, C_TYA, M_IMP, K_N); set(0xB8, C_CLV, M_IMP, K_N); set(0xD8, C_CLD, M_IMP, K_N); set(0xF8, C_SED, M_IMP, K_N);
    set(0x8A, C_TXA, M_IMP, K_N); set(0x9A, C_TXS, M_IMP, K_N); set(0xAA, C_TAX, M_IMP, K_N); set(0xBA, C_TSX, M_IMP, K_N);
    set(0xCA, C_DEX, M_IMP, K_N); set(0xEA, C_NOP, M_IMP, K_N);
}


// ---- hot decode word ---------------------------------------------------------------------------
// The kernel does not switch on 60 operation classes: they are folded into ~25 execution groups
// whose variants are selected by data bits (which register a load targets, which flag a CLC/SEC
// touches, ...), so that lanes running LDA and LDX, or CLC and SED, stay converged.
//   [2:0] am   0 none 1 imm 2 zp-family 3 abs-family 4 (zp,x) 5 (zp),y 6 (abs) 7 rel
//   [4:3] idx  0 none 1 X 2 Y          [6:5] kind (bit0 reads memory, bit1 writes memory)
//   [10:7] base cycles                 [12:11] length      [17:13] group
//   [18] page-cross penalty applies (indexed absolute READ)    [19] the reads are stack pulls
//   [21:20] number of bus reads before execute (0, 1 operand, or 1..3 pulls)           [31:22] data
enum Grp : uint32_t {
    G_NOP, G_LOAD, G_STORE, G_XFER, G_PUSH, G_PULL, G_LOGIC, G_ADC, G_SBC, G_CMP, G_BIT, G_SHIFT,
    G_INCDEC_MEM, G_INCDEC_REG, G_FLAG, G_JMP, G_JSR, G_RTS, G_RTI, G_BRK, G_BRANCH, G_RMW_ALU,
    G_ANC, G_ALR, G_SBX, G_COUNT
};
static inline uint32_t encode_hot(uint32_t e, uint32_t op)
{
    uint32_t mode = e & 15, kind = (e >> 4) & 3, cyc = (e >> 6) & 15, cls = e >> 10;
    static const uint32_t AM[13] = { 0, 1, 2, 2, 2, 3, 3, 3, 4, 5, 7, 6, 0 };
    static const uint32_t IX[13] = { 0, 0, 0, 1, 2, 0, 1, 2, 0, 2, 0, 0, 0 };
    static const uint32_t LEN[13] = { 1, 2, 2, 2, 2, 3, 3, 3, 2, 2, 2, 3, 1 };
    uint32_t grp = G_NOP, data = 0;
    switch (cls) {
    case C_LDA: grp = G_LOAD; data = 1; break;
    case C_LDX: grp = G_LOAD; data = 2; break;
    case C_LDY: grp = G_LOAD; data = 4; break;
    case C_LAX: grp = G_LOAD; data = 3; break;
    case C_STA: grp = G_STORE; data = 1; break;
    case C_STX: grp = G_STORE; data = 2; break;
    case C_STY: grp = G_STORE; data = 4; break;
    case C_SAX: grp = G_STORE; data = 3; break;
    // transfers: data = src (0 A, 1 X, 2 Y, 3 SP) | dst << 2 | setflags << 4
    case C_TAX: grp = G_XFER; data = 0 | (1 << 2) | 16; break;
    case C_TAY: grp = G_XFER; data = 0 | (2 << 2) | 16; break;
    case C_TXA: grp = G_XFER; data = 1 | (0 << 2) | 16; break;
    case C_TYA: grp = G_XFER; data = 2 | (0 << 2) | 16; break;
    case C_TSX: grp = G_XFER; data = 3 | (1 << 2) | 16; break;
    case C_TXS: grp = G_XFER; data = 1 | (3 << 2); break;
    case C_PHA: grp = G_PUSH; data = 0; break;
    case C_PHP: grp = G_PUSH; data = 1; break;
    case C_PLA: grp = G_PULL; data = 0; break;
    case C_PLP: grp = G_PULL; data = 1; break;
    case C_AND: grp = G_LOGIC; data = 0; break;
    case C_ORA: grp = G_LOGIC; data = 1; break;
    case C_EOR: grp = G_LOGIC; data = 2; break;
    case C_ADC: grp = G_ADC; break;
    case C_SBC: grp = G_SBC; break;
    case C_CMP: grp = G_CMP; data = 0; break;
    case C_CPX: grp = G_CMP; data = 1; break;
    case C_CPY: grp = G_CMP; data = 2; break;
    case C_BIT: grp = G_BIT; break;
    case C_ASL: grp = G_SHIFT; data = 0; break;
    case C_LSR: grp = G_SHIFT; data = 1; break;
    case C_ROL: grp = G_SHIFT; data = 2; break;
    case C_ROR: grp = G_SHIFT; data = 3; break;
    case C_INC: grp = G_INCDEC_MEM; data = 1; break;
    case C_DEC: grp = G_INCDEC_MEM; data = 0xFF; break;
    // data = reg (1 X, 2 Y) | delta << 2
    case C_INX: grp = G_INCDEC_REG; data = 1 | (1 << 2); break;
    case C_INY: grp = G_INCDEC_REG; data = 2 | (1 << 2); break;
    case C_DEX: grp = G_INCDEC_REG; data = 1 | (0xFF << 2); break;
    case C_DEY: grp = G_INCDEC_REG; data = 2 | (0xFF << 2); break;
    // data = flag mask | set << 8
    case C_CLC: grp = G_FLAG; data = 0x01; break;
    case C_SEC: grp = G_FLAG; data = 0x01 | 256; break;
    case C_CLI: grp = G_FLAG; data = 0x04; break;
    case C_SEI: grp = G_FLAG; data = 0x04 | 256; break;
    case C_CLV: grp = G_FLAG; data = 0x40; break;
    case C_CLD: grp = G_FLAG; data = 0x08; break;
    case C_SED: grp = G_FLAG; data = 0x08 | 256; break;
    case C_JMP: grp = G_JMP; break;
    case C_JSR: grp = G_JSR; break;
    case C_RTS: grp = G_RTS; break;
    case C_RTI: grp = G_RTI; break;
    case C_BRK: grp = G_BRK; break;
    // data = flag select (0 N, 1 V, 2 C, 3 Z) | wanted value << 2
    case C_BRA: grp = G_BRANCH; data = (op >> 6) | (((op >> 5) & 1) << 2); break;
    case C_SLO: grp = G_RMW_ALU; data = 0; break;
    case C_RLA: grp = G_RMW_ALU; data = 1; break;
    case C_SRE: grp = G_RMW_ALU; data = 2; break;
    case C_RRA: grp = G_RMW_ALU; data = 3; break;
    case C_DCP: grp = G_RMW_ALU; data = 4; break;
    case C_ISC: grp = G_RMW_ALU; data = 5; break;
    case C_ANC: grp = G_ANC; break;
    case C_ALR: grp = G_ALR; break;
    case C_SBX: grp = G_SBX; break;
    default: grp = G_NOP; break;
    }
    if (mode == M_ACC) kind = K_N;
    const uint32_t npull = grp == G_PULL ? 1u : (grp == G_RTS ? 2u : (grp == G_RTI ? 3u : 0u));
    const uint32_t nrd = npull ? npull : (((kind & 1) && mode != M_IMM) ? 1u : 0u);
    const uint32_t pen = (kind == K_R && (mode == M_ABX || mode == M_ABY)) ? 1u : 0u;
    return AM[mode] | (IX[mode] << 3) | (kind << 5) | (cyc << 7) | (LEN[mode] << 11) | (grp << 13) | (pen << 18) |
           ((npull ? 1u : 0u) << 19) | (nrd << 20) | (data << 22);
}
// the table the kernel uses
static inline void build_hot_table(uint32_t *t)
{
    uint32_t e[256];
    build_decode_table(e);
    for (uint32_t op = 0; op < 256; op++) t[op] = encode_hot(e[op], op);
}

// predecoded cartridge: entry[off] = { hot word of rom[off], rom[off+1] | rom[off+2] << 8 } with the
// operand offsets wrapped inside the bank window (window = mask + 1 bytes), as the bus would see them
static inline void build_predecode(const uint8_t *rom, uint32_t rom_size, uint32_t mask, uint32_t *out /* 2 * rom_size */)
{
    uint32_t hot[256];
    build_hot_table(hot);
    for (uint32_t off = 0; off < rom_size; off++) {
        uint32_t win = off & ~mask;
        uint32_t o1 = win | ((off + 1) & mask), o2 = win | ((off + 2) & mask);
        out[2 * off] = hot[rom[off]];
        out[2 * off + 1] = rom[o1] | ((uint32_t)rom[o2] << 8);
    }
}

// translated cartridges (csrc/gen/xl_*.cuh): write case number k+1 of offsets[k] into bits 16..31 of the operand word
static inline void stamp_cases(uint32_t *pre, const uint16_t *offsets, uint32_t n)
{
    for (uint32_t k = 0; k < n; k++) pre[2 * offsets[k] + 1] |= (k + 1) << 16;
}

// identifies a cartridge image (FNV-1a, as tools/xlate_rom.cpp): a translated core belongs to exactly one image
static inline uint32_t rom_hash(const uint8_t *rom, uint32_t n)
{
    uint32_t h = 2166136261u;
    for (uint32_t i = 0; i < n; i++) { h ^= rom[i]; h *= 16777619u; }
    return h;
}

// ---- game constants handed to the kernel ---------------------------------------------------------
struct GameParams {
    uint32_t rom_mask;        // byte mask inside one bank window (0x7FF / 0xFFF)
    uint32_t bank_scheme, controller, n_actions;
    uint8_t  actions[20];
    uint8_t  score_addr[4], score2_addr[4];
    uint8_t  lives_addr, lives_shift, lives_mask, term_addr, term_mask, term_value, pad0, pad1;
};

enum : uint32_t { FC = 1, FZ = 2, FI = 4, FD = 8, FB = 16, FU = 32, FV = 64, FN = 128 };

// byte offsets of blob fields (checked against the struct in stm_kernels.cu)
enum : uint32_t {
    O_PC = 128, O_A = 130, O_X = 131, O_Y = 132, O_SP = 133, O_P = 134, O_BANK = 135, O_LINE = 136,
    O_HCYC = 138, O_RX = 139, O_FRAME = 140, O_TIMER = 144, O_TSHIFT = 148, O_TFLAG = 149,
    O_SWCHA_OUT = 150, O_SWACNT = 151, O_SWCHB_OUT = 152, O_SWBCNT = 153, O_ACTION = 154, O_FLAGS = 155,
    O_CX = 156, O_CHARGE = 158, O_INPT_LATCH = 159, O_PADDLE = 160, O_POS = 162,
    O_VSYNC = 168, O_VBLANK = 169, O_NUSIZ0 = 170, O_NUSIZ1 = 171, O_COLUP0 = 172, O_COLUP1 = 173,
    O_COLUPF = 174, O_COLUBK = 175, O_CTRLPF = 176, O_REFP0 = 177, O_REFP1 = 178, O_PF0 = 179,
    O_PF1 = 180, O_PF2 = 181, O_GRP0 = 182, O_GRP1 = 183, O_GRP0_OLD = 184, O_GRP1_OLD = 185,
    O_ENAM0 = 186, O_ENAM1 = 187, O_ENABL = 188, O_ENABL_OLD = 189, O_HMP0 = 190, O_HMP1 = 191,
    O_HMM0 = 192, O_HMM1 = 193, O_HMBL = 194, O_VDELP0 = 195, O_VDELP1 = 196, O_VDELBL = 197,
    O_RESMP0 = 198, O_RESMP1 = 199, O_SCORE_PREV = 200
};

STM_FN uint32_t brev8(uint32_t v)
{
#ifdef __CUDA_ARCH__
    return __brev(v) >> 24;
#else
    v = ((v & 0xF0) >> 4) | ((v & 0x0F) << 4);
    v = ((v & 0xCC) >> 2) | ((v & 0x33) << 2);
    v = ((v & 0xAA) >> 1) | ((v & 0x55) << 1);
    return v;
#endif
}
STM_FN uint32_t brev32(uint32_t v)
{
#ifdef __CUDA_ARCH__
    return __brev(v);
#else
    v = ((v >> 1) & 0x55555555u) | ((v & 0x55555555u) << 1);
    v = ((v >> 2) & 0x33333333u) | ((v & 0x33333333u) << 2);
    v = ((v >> 4) & 0x0F0F0F0Fu) | ((v & 0x0F0F0F0Fu) << 4);
    v = ((v >> 8) & 0x00FF00FFu) | ((v & 0x00FF00FFu) << 8);
    return (v >> 16) | (v << 16);
#endif
}
STM_FN uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh)   // sh in {0,8,16,24}
{
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, sh);
#else
    return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}
// each of the low 8 bits replicated 4x: bit i -> bits 4i..4i+3
STM_FN uint32_t spread8x4(uint32_t b)
{
    b = (b | (b << 12)) & 0x000F000Fu;
    b = (b | (b << 6))  & 0x03030303u;
    b = (b | (b << 3))  & 0x11111111u;
    return b * 0xFu;
}
// low 4 bits -> 4 byte masks 0x00 / 0xFF
STM_FN uint32_t nib_to_bytes(uint32_t n)
{
    return (((n & 0xF) * 0x00204081u) & 0x01010101u) * 0xFFu;
}
STM_FN uint32_t sel_bytes(uint32_t under, uint32_t over, uint32_t m) { return (under & ~m) | (over & m); }


// Rows of the 210-row native frame that the 84x84 bilinear resize never samples (row % 5 == 2)
// are not drawn at all; nothing downstream reads them.
STM_FN bool row_sampled(int row) { return (uint32_t)row < (uint32_t)STM_VIS_H && (row % 5) != 2; }

// ---- TIA line renderer --------------------------------------------------------------------------
// Renders pixels [rx, xe) of the current scanline from the TIA registers in the state blob and
// latches collisions.  It is the one large piece of code in the step kernel, so it is compiled
// OUT OF LINE (a real call, arguments by value in registers): the instruction loop that calls it
// then fits the instruction cache, and its five call sites share one copy.
template <class M>
struct TiaLine {
    M m;
    uint32_t line, rx, flags;
    bool render;
    // ---------------- TIA rendering: 32-pixel words from object masks ----------------
    // bits of `pat` (width <= 32, bit 0 = leftmost pixel) placed at ring position `start` (0..159),
    // as seen by word w (pixels 32w..32w+31); the 160-pixel line is a ring.
    STM_DEV uint32_t place(uint32_t pat, int start, int w)
    {
        int r = start - 32 * w;
        uint32_t out = 0;
        if (r > -32 && r < 32) out = r >= 0 ? pat << r : pat >> (-r);
        r -= 160;
        if (r > -32 && r < 32) out |= r >= 0 ? pat << r : pat >> (-r);
        return out;
    }
    STM_DEV uint32_t copies(uint32_t pat, uint32_t pos, uint32_t nusiz, int w)
    {
        uint32_t mode = nusiz & 7;
        uint32_t out = place(pat, (int)pos, w);
        // second / third copy offsets: 1:{16} 2:{32} 3:{16,32} 4:{64} 6:{32,64}
        uint32_t o1 = (0x02041210u >> (4 * mode)) & 0xFu;      // in units of 16 px: modes 1,3 -> 1; 2,6 -> 2; 4 -> 4
        uint32_t o2 = mode == 3 ? 2u : (mode == 6 ? 4u : 0u);
        if (o1) { uint32_t s = pos + 16 * o1; if (s >= 160) s -= 160; out |= place(pat, (int)s, w); }
        if (o2) { uint32_t s = pos + 16 * o2; if (s >= 160) s -= 160; out |= place(pat, (int)s, w); }
        return out;
    }
    STM_DEV uint32_t player_pat(uint32_t grp, uint32_t nusiz, uint32_t refp)
    {
        uint32_t mode = nusiz & 7;
        uint32_t g = (refp & 8) ? grp : brev8(grp);          // bit 0 = leftmost pixel
        if (mode == 5) {                                      // double width: each bit -> 2
            g = (g | (g << 4)) & 0x0F0Fu; g = (g | (g << 2)) & 0x3333u; g = (g | (g << 1)) & 0x5555u;
            return g * 3u;
        }
        if (mode == 7) return spread8x4(g);                   // quad width
        return g;
    }

    // Rows of the 210-row native frame that the 84x84 bilinear resize never samples (row % 5 == 2)
    // are not drawn at all; nothing downstream reads them.

    // four pixels (one nibble of every layer mask) -> four packed colour bytes
    STM_DEV uint32_t compose4(uint32_t l0, uint32_t l1, uint32_t blm, uint32_t pfm, uint32_t sh, uint32_t c_p0,
                              uint32_t c_p1, uint32_t c_pf, uint32_t c_bk, uint32_t pfcol, bool pfp)
    {
        uint32_t b0 = nib_to_bytes(l0 >> sh), b1 = nib_to_bytes(l1 >> sh);
        uint32_t bb = nib_to_bytes(blm >> sh), bf = nib_to_bytes(pfm >> sh);
        uint32_t word = c_bk;
        if (pfp) {
            word = sel_bytes(word, c_p1, b1); word = sel_bytes(word, c_p0, b0);
            word = sel_bytes(word, c_pf, bf | bb);
        } else {
            word = sel_bytes(word, pfcol, bf); word = sel_bytes(word, c_pf, bb);
            word = sel_bytes(word, c_p1, b1);  word = sel_bytes(word, c_p0, b0);
        }
        return word;
    }
    // draw pixels [lo,hi) of word w (32 px).  A word holding only playfield/background (the usual
    // case) costs one select per 4-pixel column and is stored 16 pixels at a time.
    STM_DEV void emit_word(int row, int w, int lo, int hi, uint32_t seg, uint32_t l0, uint32_t l1, uint32_t blm,
                           uint32_t pfm, uint32_t c_p0, uint32_t c_p1, uint32_t c_pf, uint32_t c_bk, bool pfp,
                           bool score, bool comb)
    {
        const bool plain = (l0 | l1 | blm) == 0;
        #pragma unroll 1
        for (int q = 0; q < 2; q++) {                          // two 16-pixel quads per word
            int qlo = 16 * q, qhi = qlo + 16;
            if (hi <= qlo || lo >= qhi) continue;
            uint32_t wd[4];
            #pragma unroll
            for (int j = 0; j < 4; j++) {
                uint32_t sh = (uint32_t)(qlo + 4 * j);
                uint32_t pfcol = score ? ((32 * w + (int)sh) < 80 ? c_p0 : c_p1) : c_pf;
                uint32_t word = plain ? (((pfm >> sh) & 0xFu) ? pfcol : c_bk)
                                      : compose4(l0, l1, blm, pfm, sh, c_p0, c_p1, c_pf, c_bk, pfcol, pfp);
                if (comb && w == 0 && q == 0 && j < 2) word = 0;
                wd[j] = word;
            }
            if (((seg >> qlo) & 0xFFFFu) == 0xFFFFu) m.px16((uint32_t)row, (uint32_t)(2 * w + q), wd[0], wd[1], wd[2], wd[3]);
            else {
                #pragma unroll 1
                for (int j = 0; j < 4; j++) {
                    uint32_t nm = (seg >> (qlo + 4 * j)) & 0xFu;
                    uint32_t g = 8u * (uint32_t)w + 4u * (uint32_t)q + (uint32_t)j;
                    if (nm == 0xFu) m.px4((uint32_t)row, g, wd[j]);
                    else for (uint32_t k = 0; k < 4; k++)
                        if (nm & (1u << k)) m.px1((uint32_t)row, 4u * g + k, (wd[j] >> (8u * k)) & 0xFFu);
                }
            }
        }
    }
    // a span of pixels drawn in one colour pattern (used for blanked lines)
    STM_DEV void emit_span(int row, uint32_t x0, uint32_t x1, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t,
                           uint32_t, uint32_t, uint32_t, bool, bool, bool, bool)
    {
        int w0 = (int)(x0 >> 5), w1 = (int)((x1 - 1) >> 5);
        #pragma unroll 1
        for (int w = w0; w <= w1; w++) {
            int lo = (int)x0 - 32 * w; if (lo < 0) lo = 0;
            int hi = (int)x1 - 32 * w; if (hi > 32) hi = 32;
            uint32_t seg = (hi == 32 ? 0xFFFFFFFFu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
            emit_word(row, w, lo, hi, seg, 0, 0, 0, 0, 0, 0, 0, 0, false, false, false);
        }
    }

    STM_DEV void catchup(uint32_t xe)
    {
        int row = (int)line - STM_YSTART;
        bool draw = render && row_sampled(row);
        uint32_t vblank = m.ld8(O_VBLANK);
        bool vb = (vblank & 2) != 0;
        uint32_t g0 = (m.ld8(O_VDELP0) & 1) ? m.ld8(O_GRP0_OLD) : m.ld8(O_GRP0);
        uint32_t g1 = (m.ld8(O_VDELP1) & 1) ? m.ld8(O_GRP1_OLD) : m.ld8(O_GRP1);
        uint32_t enbl = (m.ld8(O_VDELBL) & 1) ? m.ld8(O_ENABL_OLD) : m.ld8(O_ENABL);
        bool m0on = (m.ld8(O_ENAM0) & 2) && !(m.ld8(O_RESMP0) & 2);
        bool m1on = (m.ld8(O_ENAM1) & 2) && !(m.ld8(O_RESMP1) & 2);
        bool blon = (enbl & 2) != 0;
        bool movable = g0 || g1 || m0on || m1on || blon;
        if (!draw && (vb || !movable)) { rx = xe; return; }

        int w0 = (int)(rx >> 5), w1 = (int)((xe - 1) >> 5);
        if (draw && vb) {                                    // blanked line: black pixels, no collisions
            emit_span(row, rx, xe, 0, 0, 0, 0, 0, 0, 0, 0, false, false, false, true);
            rx = xe;
            return;
        }

        uint32_t ctrlpf = m.ld8(O_CTRLPF);
        uint32_t nusiz0 = m.ld8(O_NUSIZ0), nusiz1 = m.ld8(O_NUSIZ1);
        // playfield as 40 four-pixel columns: low word = columns 0..31, high = 32..39
        uint32_t pf20 = (m.ld8(O_PF0) >> 4) | (brev8(m.ld8(O_PF1)) << 4) | (m.ld8(O_PF2) << 12);
        uint32_t pfr = (ctrlpf & 1) ? (brev32(pf20) >> 12) : pf20;
        uint32_t pf_lo = pf20 | (pfr << 20), pf_hi = pfr >> 12;
        uint32_t pat0 = g0 ? player_pat(g0, nusiz0, m.ld8(O_REFP0)) : 0;
        uint32_t pat1 = g1 ? player_pat(g1, nusiz1, m.ld8(O_REFP1)) : 0;
        uint32_t mp0 = (1u << (1u << ((nusiz0 >> 4) & 3))) - 1u;
        uint32_t mp1 = (1u << (1u << ((nusiz1 >> 4) & 3))) - 1u;
        uint32_t bp  = (1u << (1u << ((ctrlpf >> 4) & 3))) - 1u;
        uint32_t pos_p0 = m.ld8(O_POS + 0), pos_p1 = m.ld8(O_POS + 1), pos_m0 = m.ld8(O_POS + 2),
                 pos_m1 = m.ld8(O_POS + 3), pos_bl = m.ld8(O_POS + 4);
        uint32_t cx = 0;
        uint32_t c_p0 = 0, c_p1 = 0, c_pf = 0, c_bk = 0;
        bool comb = (flags & STM_F_HMOVE_BLANK) != 0;
        bool pfp = (ctrlpf & 4) != 0, score = (ctrlpf & 2) != 0 && !pfp;
        if (draw) {
            // pixels are stored as GREY values (the palette look-up happens here, four times per line,
            // not per pixel in the frame kernel); colour index 0 is grey 0, so blank rows stay zero
            c_p0 = m.grey(m.ld8(O_COLUP0) >> 1) * 0x01010101u; c_p1 = m.grey(m.ld8(O_COLUP1) >> 1) * 0x01010101u;
            c_pf = m.grey(m.ld8(O_COLUPF) >> 1) * 0x01010101u; c_bk = m.grey(m.ld8(O_COLUBK) >> 1) * 0x01010101u;
        }
        #pragma unroll 1
        for (int w = w0; w <= w1; w++) {
            int lo = (int)rx - 32 * w; if (lo < 0) lo = 0;
            int hi = (int)xe - 32 * w; if (hi > 32) hi = 32;
            uint32_t seg = (hi == 32 ? 0xFFFFFFFFu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
            uint32_t pfcols = w < 4 ? (pf_lo >> (8 * w)) & 0xFFu : pf_hi & 0xFFu;
            uint32_t pfm = spread8x4(pfcols) & seg;
            uint32_t l0 = 0, l1 = 0, blm = 0;
            if (movable) {
                uint32_t p0m = pat0 ? copies(pat0, pos_p0, nusiz0, w) & seg : 0;
                uint32_t p1m = pat1 ? copies(pat1, pos_p1, nusiz1, w) & seg : 0;
                uint32_t m0m = m0on ? copies(mp0, pos_m0, nusiz0, w) & seg : 0;
                uint32_t m1m = m1on ? copies(mp1, pos_m1, nusiz1, w) & seg : 0;
                blm = blon ? place(bp, (int)pos_bl, w) & seg : 0;
                l0 = p0m | m0m; l1 = p1m | m1m;
                if (!vb && (l0 | l1 | blm)) {
                    if (m0m & p0m) cx |= 1u << 0;
                    if (m0m & p1m) cx |= 1u << 1;
                    if (m1m & p1m) cx |= 1u << 2;
                    if (m1m & p0m) cx |= 1u << 3;
                    if (p0m & blm) cx |= 1u << 4;
                    if (p0m & pfm) cx |= 1u << 5;
                    if (p1m & blm) cx |= 1u << 6;
                    if (p1m & pfm) cx |= 1u << 7;
                    if (m0m & blm) cx |= 1u << 8;
                    if (m0m & pfm) cx |= 1u << 9;
                    if (m1m & blm) cx |= 1u << 10;
                    if (m1m & pfm) cx |= 1u << 11;
                    if (blm & pfm) cx |= 1u << 13;
                    if (m0m & m1m) cx |= 1u << 14;
                    if (p0m & p1m) cx |= 1u << 15;
                }
            }
            if (draw)
                emit_word(row, w, lo, hi, seg, l0, l1, blm, pfm, c_p0, c_p1, c_pf, c_bk, pfp, score, comb);
        }
        if (cx) m.st16(O_CX, m.ld16(O_CX) | cx);
        rx = xe;
    }

};

template <class M>
STM_NOINLINE uint32_t tia_catchup(M m, uint32_t line, uint32_t rx, uint32_t xe, uint32_t flags, bool render)
{
    TiaLine<M> t{m, line, rx, flags, render};
    t.catchup(xe);
    return t.rx;
}

// decimal-mode ADC: returns a | p << 8 | zn << 16 (p: C, I, D, V bits as in Emu::p)
STM_NOINLINE uint32_t adc_decimal(uint32_t a, uint32_t v, uint32_t p)
{
    uint32_t cin = p & FC;
    uint32_t al = (a & 0x0F) + (v & 0x0F) + cin;
    if (al > 9) al += 6;
    uint32_t ah = (a >> 4) + (v >> 4) + (al > 0x0F ? 1u : 0u);
    uint32_t bin = (a + v + cin) & 0xFF;
    p &= ~(uint32_t)(FV | FC);
    uint32_t zn = (bin ? 1u : 0u) | ((ah & 8) << 5);           // Z from the binary sum, N from the high digit
    if (~(a ^ v) & (a ^ (ah << 4)) & 0x80) p |= FV;
    if (ah > 9) ah += 6;
    if (ah > 15) p |= FC;
    a = ((ah << 4) | (al & 0x0F)) & 0xFF;
    return a | (p << 8) | (zn << 16);
}
// decimal-mode SBC: the accumulator result (flags are the binary ones)
STM_NOINLINE uint32_t sbc_decimal(uint32_t a, uint32_t v, uint32_t bc)
{
    int al = (int)(a & 0x0F) - (int)(v & 0x0F) - (int)bc;
    int ah = (int)(a >> 4) - (int)(v >> 4);
    if (al & 0x10) { al -= 6; ah--; }
    if (ah & 0x10) ah -= 6;
    return (uint32_t)(((ah << 4) | (al & 0x0F)) & 0xFF);
}

// native row `src` of the frame being drawn -> row `dst` (whole-line reuse, see Emu::end_line)
template <class M>
STM_NOINLINE void tia_copy_row(M m, uint32_t dst, uint32_t src) { m.copy_row(dst, src); }   // (host test aid only; the device links rows)

// ---- the per-env machine ---------------------------------------------------------------------------
// M must provide:
//   uint32_t ld8(off) / void st8(off, v) / ld16 / st16 / ld32 / st32     blob access (off = byte offset)
//   uint32_t rom8(idx)                                                     cartridge image
//   void     fetch(off, &w, &ops)   predecoded instruction at image offset off: w = hot decode word of
//                                   the opcode there, ops = operand bytes b1 | b2 << 8 (build_predecode)
//   uint32_t opinfo(op)                                                    hot decode word (build_hot_table)
//   uint32_t grey(idx)        NTSC colour index (COLUxx >> 1) -> 8-bit grey (include/stm_palette.h)
//   void px16(row, q, w0..w3) store 16 pixels (grey values) of native row `row`, quad q (0..9): four packed words
//   void px4(row, g, word)    store 4 pixels, group g (0..39)
//   void px1(row, x, v)       store 1 pixel
//   void open_row(row)        pixels of native row `row` are about to be stored (called before every catch-up that draws)
//   void link_row(dst, src)   native row dst shows exactly what row src shows (whole-line reuse); src is the row opened last
//   void zero_row(row)        native row `row` is black (rows the frame never reached)
//
// One instruction = fetch -> decode -> address -> read phase -> execute (register-only switch over
// ~25 groups) -> write phase -> retire.  Every memory access goes through ONE read site and ONE
// write site, so the TIA/RIOT slow paths (and the big catch-up renderer behind them) exist once in
// the kernel instead of once per opcode; stack pulls/pushes are sequenced through the same sites.
// The instruction loop of one scanline, as a policy of Emu::run_frame.  InterpCore interprets: fetch,
// decode, dispatch per instruction.  (Lowest-program-counter-first lane scheduling was measured neutral
// for the interpreter on B200; the translated cores use it, see tools/xlate_rom.cpp.)
struct InterpCore {
    template <class E>
    static STM_DEV void run_line(E &e, uint32_t lanes, bool over)
    {
        (void)lanes;
        if (!over) while (!e.exec_core()) { }
    }
};

template <class M>
struct Emu {
    M &m;
    const GameParams &gp;
    // hot state (registers)
    uint32_t pc, a, x, y, sp;
    uint32_t p;             // C, I, D, V only (mask 0x4D); N and Z live in zn
    uint32_t zn;            // Z <=> (zn & 0xFF) == 0 ; N <=> (zn & 0x180) != 0
    int32_t  hcyc;          // CPU cycle in the line; may be transiently negative inside an instruction
    int32_t  timer;
    uint32_t bank;
    // warm state
    uint32_t line, rx, flags, end_line_no;
    bool     render;
    // instruction-loop break request, added to hcyc in the loop test: WSYNC strobed / frame over
    int32_t  brk;
    static constexpr int32_t BRK_WSYNC = 0x100, BRK_DONE = 0x200;

    STM_DEV Emu(M &mm, const GameParams &g) : m(mm), gp(g) {}

    STM_DEV uint32_t get_p() const
    { return (p & 0x4Du) | FU | ((zn & 0x180u) ? (uint32_t)FN : 0u) | ((zn & 0xFFu) ? 0u : (uint32_t)FZ); }
    STM_DEV void set_p(uint32_t v) { p = v & 0x4Du; zn = ((v & FN) << 1) | ((v & FZ) ? 0u : 1u); }

    STM_DEV void load()
    {
        pc = m.ld16(O_PC); a = m.ld8(O_A); x = m.ld8(O_X); y = m.ld8(O_Y); sp = m.ld8(O_SP); set_p(m.ld8(O_P));
        bank = m.ld8(O_BANK); line = m.ld16(O_LINE); hcyc = (int32_t)m.ld8(O_HCYC); rx = m.ld8(O_RX);
        flags = m.ld8(O_FLAGS); timer = (int32_t)m.ld32(O_TIMER);
        render = false; end_line_no = 0; brk = 0;
    }
    STM_DEV void store()
    {
        if (timer < -(1 << 30)) timer += (1 << 30);
        m.st16(O_PC, pc); m.st8(O_A, a); m.st8(O_X, x); m.st8(O_Y, y); m.st8(O_SP, sp);
        m.st8(O_P, get_p());
        m.st8(O_BANK, bank); m.st16(O_LINE, line); m.st8(O_HCYC, (uint32_t)hcyc); m.st8(O_RX, rx);
        m.st8(O_FLAGS, flags & 0xFFu); m.st32(O_TIMER, (uint32_t)timer);
        m.st8(O_TFLAG, timer < 0 ? 0x80u : 0u);      // the underflow latch always equals (timer < 0) between instructions
    }

    // ---------------- input ----------------
    STM_DEV bool fire_pressed() { uint32_t act = m.ld8(O_ACTION); return act == 1 || act >= 10; }
    STM_DEV uint32_t dir_bits()    // bit0 up, bit1 down, bit2 left, bit3 right
    {
        uint32_t act = m.ld8(O_ACTION);
        if (act < 2 || act > 17) return 0;
        uint32_t b = act >= 10 ? act - 8 : act;           // FIRE variants repeat codes 2..9
        // nibble table for b = 2..9: 1 8 4 2 9 5 10 6
        return (0x6A592481u >> (4 * (b - 2))) & 0xFu;
    }
    STM_DEV void update_latch()
    {
        if (!(m.ld8(O_VBLANK) & 0x40)) { m.st8(O_INPT_LATCH, 0); return; }
        if (gp.controller == STM_CTRL_JOYSTICK && fire_pressed()) m.st8(O_INPT_LATCH, m.ld8(O_INPT_LATCH) | 1);
    }
    STM_DEV void set_action(uint32_t ale)
    {
        m.st8(O_ACTION, ale);
        if (gp.controller == STM_CTRL_PADDLE) {
            uint32_t d = dir_bits();
            int pp = (int)m.ld8(O_PADDLE);
            if (d & 8) pp -= STM_PADDLE_DELTA;
            if (d & 4) pp += STM_PADDLE_DELTA;
            pp = pp < STM_PADDLE_MIN ? STM_PADDLE_MIN : (pp > STM_PADDLE_MAX ? STM_PADDLE_MAX : pp);
            m.st8(O_PADDLE, (uint32_t)pp);
        }
        update_latch();
    }

    // the row the current line draws into, announced to the memory policy before the (out-of-line, by-value) renderer runs
    STM_DEV void open_row()
    {
        const int row = (int)line - STM_YSTART;
        if (render && row_sampled(row)) m.open_row((uint32_t)row);
    }
    // ---------------- TIA catch-up (out of line: see TiaLine below) ----------------
    STM_DEV void catchup(uint32_t xe)
    {
        if (xe > STM_VIS_W) xe = STM_VIS_W;
        if (xe > rx) { open_row(); rx = tia_catchup<M>(m, line, rx, xe, flags, render); }
    }

    // Whole-line reuse.  Most scanlines of a frame are drawn under exactly the TIA state of the line
    // above (a wall, a brick band, empty field).  `flags` carries two transient bits above its stored
    // byte: F_CXCLEAN = a whole line has been processed under the current state (its collisions are
    // latched; latching them again changes nothing), F_PIXCLEAN = additionally native row
    // (flags >> 16) & 0xFF of this frame holds that line's pixels.  Every write that changes a register
    // the renderer reads clears both (tia_dirty); a line wrapped while they are set is a 160-byte row
    // copy (or nothing at all on rows/frames that are not drawn) instead of a render.
    static constexpr uint32_t F_PIXCLEAN = 0x100u, F_CXCLEAN = 0x200u;
    STM_DEV void tia_dirty() { flags &= ~(F_PIXCLEAN | F_CXCLEAN); }
    STM_DEV void end_line()
    {
        if (rx == 0) {
            int row = (int)line - STM_YSTART;
            bool draw = render && row_sampled(row);
            if (flags & (draw ? F_PIXCLEAN : F_CXCLEAN)) {
                if (draw) m.link_row((uint32_t)row, (flags >> 16) & 0xFFu);
            } else {
                open_row();
                tia_catchup<M>(m, line, 0, STM_VIS_W, flags, render);
                if (!(flags & STM_F_HMOVE_BLANK)) {
                    flags |= F_CXCLEAN;
                    if (draw) flags = (flags & 0xFF00FFFFu) | F_PIXCLEAN | ((uint32_t)row << 16);
                }
            }
        } else {
            catchup(STM_VIS_W);
            tia_dirty();
        }
        line++;
        rx = 0;
        flags &= ~(uint32_t)STM_F_HMOVE_BLANK;
        if (!(m.ld8(O_VBLANK) & 0x80)) { uint32_t c = m.ld8(O_CHARGE); if (c < 255) m.st8(O_CHARGE, c + 1); }
        if (line >= STM_MAX_LINES) { flags |= STM_F_FRAME_DONE; brk |= BRK_DONE; end_line_no = STM_MAX_LINES; line = 0; }
        if (timer < -(1 << 30)) timer += (1 << 30);
    }
    // cycle (within the line) at which the current instruction touches the bus; runs the line wrap
    // first when that cycle lies beyond the end of the line
    // kNoWrap: the caller has already checked that the access stays inside the line (translated cores send
    // the rare straddling instruction to the interpreter, so the line wrap is not inlined into every case)
    template <bool kNoWrap = false>
    STM_DEV uint32_t access_cycle(uint32_t cyc)
    {
        int32_t acc = hcyc + (int32_t)cyc - 1;
        if (!kNoWrap && STM_UNLIKELY(acc >= STM_CYC_PER_LINE)) { end_line(); hcyc -= STM_CYC_PER_LINE; acc -= STM_CYC_PER_LINE; }
        return (uint32_t)acc;
    }
    STM_DEV uint32_t x_of(uint32_t acc) { int v = (int)acc * 3 - STM_HBLANK_CLK; return v < 0 ? 0u : (uint32_t)v; }

    // ---------------- TIA registers ----------------
    STM_DEV uint32_t tia_read(uint32_t reg, uint32_t acc)
    {
        reg &= 0x0F;
        if (reg < 8) { catchup(x_of(acc)); return ((m.ld16(O_CX) >> (2 * reg)) & 3u) << 6; }
        if (reg < 12) {
            uint32_t i = reg - 8;
            if (gp.controller != STM_CTRL_PADDLE || i > 1) return 0;
            if (m.ld8(O_VBLANK) & 0x80) return 0;
            return m.ld8(O_CHARGE) >= m.ld8(O_PADDLE + i) ? 0x80u : 0u;
        }
        if (reg == 12) {
            if (gp.controller != STM_CTRL_JOYSTICK) return 0x80;
            if (fire_pressed()) return 0;
            if ((m.ld8(O_VBLANK) & 0x40) && (m.ld8(O_INPT_LATCH) & 1)) return 0;
            return 0x80;
        }
        return reg == 13 ? 0x80u : 0u;
    }
    STM_DEV void set_pos(uint32_t obj, bool player, uint32_t acc)
    {
        uint32_t clk = acc * 3;
        uint32_t v = clk < STM_HBLANK_CLK ? (player ? 3u : 2u) : (clk - STM_HBLANK_CLK + (player ? 5u : 4u)) % 160u;
        m.st8(O_POS + obj, v);
    }
    STM_DEV void hmove_obj(uint32_t obj, uint32_t hm)
    {
        int mv = (int)(int8_t)(uint8_t)hm >> 4;
        int v = (int)m.ld8(O_POS + obj) - mv + 160;
        if (v >= 160) v -= 160;
        if (v >= 160) v -= 160;
        m.st8(O_POS + obj, (uint32_t)v);
    }
    STM_DEV void tia_write(uint32_t reg, uint32_t v, uint32_t acc)
    {
        reg &= 0x3F;
        if (reg > 0x2C) return;
        if (reg == 0x02) { brk |= BRK_WSYNC; return; }
        if (reg == 0x03 || (reg >= 0x15 && reg <= 0x1A)) return;
        v &= 0xFFu;
        // Plain registers: a write of the value already there changes nothing (no catch-up needed, the
        // line stays reusable); HMxx / HMCLR are not read by the renderer at all.
        if (reg >= 0x04 && reg <= 0x0F) {                                   // NUSIZ0..PF2 are laid out in register order
            uint32_t o = O_NUSIZ0 + (reg - 4);
            if (STM_LIKELY(m.ld8(o) == v)) return;
            catchup(x_of(acc)); m.st8(o, v); tia_dirty(); return;
        }
        if (reg >= 0x1D && reg <= 0x1F) {                                   // ENAM0, ENAM1, ENABL
            uint32_t o = O_ENAM0 + (reg - 0x1D);
            if (STM_LIKELY(m.ld8(o) == v)) return;
            catchup(x_of(acc)); m.st8(o, v); tia_dirty(); return;
        }
        if (reg >= 0x20 && reg <= 0x24) { m.st8(O_HMP0 + (reg - 0x20), v); return; }   // HMP0..HMBL
        if (reg >= 0x25 && reg <= 0x27) {                                   // VDELP0, VDELP1, VDELBL follow them
            uint32_t o = O_HMP0 + (reg - 0x20);
            if (STM_LIKELY(m.ld8(o) == v)) return;
            catchup(x_of(acc)); m.st8(o, v); tia_dirty(); return;
        }
        if (reg == 0x2B) { m.st8(O_HMP0, 0); m.st8(O_HMP1, 0); m.st8(O_HMM0, 0); m.st8(O_HMM1, 0); m.st8(O_HMBL, 0); return; }
        if (reg == 0x1B) {
            uint32_t g1 = m.ld8(O_GRP1);
            if (STM_LIKELY(m.ld8(O_GRP0) == v && m.ld8(O_GRP1_OLD) == g1)) return;
            catchup(x_of(acc)); m.st8(O_GRP0, v); m.st8(O_GRP1_OLD, g1); tia_dirty(); return;
        }
        if (reg == 0x1C) {
            uint32_t g0 = m.ld8(O_GRP0), bl = m.ld8(O_ENABL);
            if (STM_LIKELY(m.ld8(O_GRP1) == v && m.ld8(O_GRP0_OLD) == g0 && m.ld8(O_ENABL_OLD) == bl)) return;
            catchup(x_of(acc)); m.st8(O_GRP1, v); m.st8(O_GRP0_OLD, g0); m.st8(O_ENABL_OLD, bl); tia_dirty(); return;
        }
        catchup(x_of(acc));
        if (reg >= 0x10 && reg <= 0x14) { set_pos(reg - 0x10, reg < 0x12, acc); tia_dirty(); return; }   // RESP0 RESP1 RESM0 RESM1 RESBL = pos[0..4]
        switch (reg) {
        case 0x00:
            if ((v & 2) && !(m.ld8(O_VSYNC) & 2)) { flags |= STM_F_FRAME_DONE; brk |= BRK_DONE; end_line_no = line; line = 0; }
            m.st8(O_VSYNC, v); break;
        case 0x01:
            if (v & 0x80) m.st8(O_CHARGE, 0);
            if (m.ld8(O_VBLANK) != v) tia_dirty();
            m.st8(O_VBLANK, v); update_latch(); break;
        case 0x28:
            if ((m.ld8(O_RESMP0) & 2) && !(v & 2)) m.st8(O_POS + STM_M0, (m.ld8(O_POS + STM_P0) + 4u) % 160u);
            m.st8(O_RESMP0, v); tia_dirty(); break;
        case 0x29:
            if ((m.ld8(O_RESMP1) & 2) && !(v & 2)) m.st8(O_POS + STM_M1, (m.ld8(O_POS + STM_P1) + 4u) % 160u);
            m.st8(O_RESMP1, v); tia_dirty(); break;
        case 0x2A:
            hmove_obj(STM_P0, m.ld8(O_HMP0)); hmove_obj(STM_P1, m.ld8(O_HMP1));
            hmove_obj(STM_M0, m.ld8(O_HMM0)); hmove_obj(STM_M1, m.ld8(O_HMM1));
            hmove_obj(STM_BL, m.ld8(O_HMBL));
            if (acc * 3 < STM_HBLANK_CLK) flags |= STM_F_HMOVE_BLANK;
            tia_dirty();
            break;
        case 0x2C: m.st16(O_CX, 0); tia_dirty(); break;
        default: break;
        }
    }

    // ---------------- RIOT ----------------
    STM_DEV uint32_t swcha_in()
    {
        if (gp.controller == STM_CTRL_PADDLE) return fire_pressed() ? 0x7Fu : 0xFFu;
        uint32_t d = dir_bits();
        // bit7 right, bit6 left, bit5 down, bit4 up; 0 = pushed
        return 0xFFu & ~(((d & 8) << 4) | ((d & 4) << 4) | ((d & 2) << 4) | ((d & 1) << 4));
    }
    // Spin-wait fast-forward.  `LDA INTIM ; BNE *-3` (AD 84 02 D0 FB) is how every 2600 program waits
    // for the end of VBLANK/overscan; run literally it is ~30% of all emulated instructions.  When the
    // read is about to return non-zero, whole iterations of the loop (which change nothing but time)
    // are skipped: time moves forward line by line exactly as the iterations would have moved it,
    // and the instruction then completes as the first iteration that could read zero.
    STM_DEV void spin_skip(uint32_t cyc, uint32_t sh)
    {
        if (sh < 3) return;                                   // the zero window [0, 2^sh) must be >= one iteration
        if (flags & STM_F_FRAME_DONE) return;                 // the frame ended inside this very instruction
        if (peek(pc + 3) != 0xD0 || peek(pc + 4) != 0xFB) return;
        uint32_t per = 7u + ((((pc + 5) ^ pc) & 0xFF00u) ? 1u : 0u);
        int32_t t_acc = timer - (int32_t)(cyc - 1);
        int32_t win = (1 << sh) - 1;
        if (t_acc <= win) return;
        int32_t j = (t_acc - win + (int32_t)per - 1) / (int32_t)per;
        int32_t lines_left = (int32_t)STM_MAX_LINES - 2 - (int32_t)line;
        if (lines_left <= 0) return;
        int32_t jmax = (lines_left * STM_CYC_PER_LINE) / (int32_t)per;
        if (j > jmax) j = jmax;
        int32_t skip = j * (int32_t)per;
        timer -= skip;
        hcyc += skip;
        while (hcyc + (int32_t)cyc - 1 >= STM_CYC_PER_LINE) { end_line(); hcyc -= STM_CYC_PER_LINE; }
    }
    STM_DEV uint32_t riot_read(uint32_t addr, uint32_t cyc, uint32_t opword)
    {
        if (addr & 4) {
            if (addr & 1) return (timer - (int32_t)(cyc - 1)) < 0 ? 0x80u : 0u;     // INSTAT
            uint32_t sh = m.ld8(O_TSHIFT);
            // LDA abs (no index) only: group LOAD, am abs, idx none, 4 cycles
            if (STM_UNLIKELY((opword & 0x3E01Fu) == ((G_LOAD << 13) | 3u) && cyc == 4)) spin_skip(cyc, sh);
            int32_t t = timer - (int32_t)(cyc - 1);
            return t >= 0 ? ((uint32_t)t >> sh) & 0xFFu : (uint32_t)t & 0xFFu;
        }
        uint32_t sel = addr & 3;
        if (sel == 0) { uint32_t d = m.ld8(O_SWACNT); return (m.ld8(O_SWCHA_OUT) & d) | (swcha_in() & ~d & 0xFFu); }
        if (sel == 1) return m.ld8(O_SWACNT);
        if (sel == 2) { uint32_t d = m.ld8(O_SWBCNT); return (m.ld8(O_SWCHB_OUT) & d) | (0x0Bu & ~d); }
        return m.ld8(O_SWBCNT);
    }
    STM_DEV void riot_write(uint32_t addr, uint32_t v, uint32_t cyc)
    {
        if (addr & 4) {
            if (addr & 0x10) {
                uint32_t sh = (0x0A060300u >> (8 * (addr & 3))) & 0xFFu;
                m.st8(O_TSHIFT, sh);
                // the write happens at cycle cyc-1 of the instruction; retire subtracts all `cyc` cycles
                timer = (int32_t)(v << sh) - 1 + (int32_t)(cyc - 1);
            }
            return;
        }
        m.st8(O_SWCHA_OUT + (addr & 3), v);     // SWCHA_OUT, SWACNT, SWCHB_OUT, SWBCNT are consecutive
    }

    // ---------------- bus ----------------
    STM_DEV uint32_t rom_at(uint32_t addr) { return m.rom8((bank << 12) | (addr & gp.rom_mask)); }
    STM_DEV void hotspot(uint32_t addr)
    {
        if (gp.bank_scheme == STM_BANK_F8) {
            if (addr == 0x1FF8) bank = 0; else if (addr == 0x1FF9) bank = 1;
        }
    }
    STM_DEV uint32_t peek(uint32_t addr)
    {
        addr &= 0x1FFF;
        if (addr & 0x1000) return rom_at(addr);
        if ((addr & 0x280) == 0x80) return m.ld8(addr & 0x7F);
        return 0;
    }
    STM_DEV uint32_t peek16zp(uint32_t z) { return peek(z & 0xFF) | (peek((z + 1) & 0xFF) << 8); }
    // the single read site / write site of the machine
    // what a translated instruction knows at compile time about the address it touches
    enum AddrCls : uint32_t { A_ANY, A_ROM, A_ZP };
    template <bool kNoWrap = false>
    STM_DEV uint32_t bus_read(uint32_t addr, uint32_t cyc, uint32_t opword, uint32_t cls = A_ANY)
    {
        addr &= 0x1FFF;
        if (cls == A_ROM) return rom_at(addr);                          // cartridge space, no hotspot in range
        if (cls == A_ZP) {                                              // zero page: RIOT RAM or a TIA read register
            if (addr & 0x80) return m.ld8(addr & 0x7F);
            return tia_read(addr, access_cycle<kNoWrap>(cyc));
        }
        if ((addr & 0x1280) == 0x80) return m.ld8(addr & 0x7F);          // RAM: the common case
        if (addr & 0x1000) { hotspot(addr); return rom_at(addr); }
        uint32_t acc = access_cycle<kNoWrap>(cyc);
        if (!(addr & 0x80)) return tia_read(addr, acc);
        return riot_read(addr, cyc, opword);
    }
    template <bool kNoWrap = false>
    STM_DEV void bus_write(uint32_t addr, uint32_t v, uint32_t cyc, uint32_t cls = A_ANY)
    {
        addr &= 0x1FFF;
        if (cls == A_ZP) {
            if (addr & 0x80) { m.st8(addr & 0x7F, v); return; }
            tia_write(addr, v, access_cycle<kNoWrap>(cyc));
            return;
        }
        if ((addr & 0x1280) == 0x80) { m.st8(addr & 0x7F, v); return; }
        if (STM_UNLIKELY(addr & 0x1000)) { hotspot(addr); return; }
        uint32_t acc = access_cycle<kNoWrap>(cyc);
        if (!(addr & 0x80)) { tia_write(addr, v, acc); return; }
        riot_write(addr, v, cyc);
    }

    // ---------------- ALU helpers (register only) ----------------
    STM_DEV void adc(uint32_t v)
    {
        if (STM_UNLIKELY(p & FD)) {
            uint32_t r = adc_decimal(a, v, p);                 // out of line: one copy in a translated kernel
            a = r & 0xFF; p = (r >> 8) & 0xFF; zn = r >> 16;
        } else {
            uint32_t cin = p & FC;
            uint32_t s = a + v + cin;
            p = (p & ~(uint32_t)(FV | FC)) | (s >> 8) | (((~(a ^ v) & (a ^ s)) >> 1) & FV);
            a = s & 0xFF;
            zn = a;
        }
    }
    STM_DEV void sbc(uint32_t v)
    {
        uint32_t bc = (p & FC) ^ 1u;
        uint32_t s = a - v - bc;
        uint32_t res = s & 0xFF;
        const bool dec = (p & FD) != 0;
        const uint32_t a0 = a;
        p = (p & ~(uint32_t)(FV | FC)) | (((s >> 8) & 1u) ^ 1u) | ((((a ^ v) & (a ^ s)) >> 1) & FV);
        zn = res;
        if (STM_UNLIKELY(dec)) res = sbc_decimal(a0, v, bc);
        a = res;
    }
    STM_DEV void cmp(uint32_t r, uint32_t v)
    {
        uint32_t d = r - v;
        p = (p & ~(uint32_t)FC) | (((d >> 8) & 1u) ^ 1u);
        zn = d & 0xFF;
    }
    // shift group: sub 0 ASL 1 LSR 2 ROL 3 ROR
    STM_DEV uint32_t shift(uint32_t v, uint32_t sub)
    {
        uint32_t cin = (sub & 2) ? (p & FC) : 0u;
        uint32_t cout, r;
        if (sub & 1) { cout = v & 1; r = (v >> 1) | (cin << 7); }
        else         { cout = v >> 7; r = ((v << 1) | cin) & 0xFF; }
        p = (p & ~(uint32_t)FC) | cout;
        zn = r;
        return r;
    }

    // ---------------- one instruction ----------------
    // Returns true when the caller must leave its instruction loop: the instruction ran past the end of
    // the scanline or strobed WSYNC (call finish_line()), or it ended the frame.  Keeping the line wrap
    // out of here lets the step kernel run all lanes of a warp to the end of their scanline and wrap —
    // i.e. render — them together.
    STM_DEV bool exec_core()
    {
        // fetch + decode.  Cartridge code is predecoded (one 64-bit record per image byte: the hot
        // decode word of the opcode there and its two operand bytes), so the common case is one load.
        uint32_t w, b1, b2;
        uint32_t a13 = pc & 0x1FFF;
        if (STM_LIKELY(a13 - 0x1000u <= 0xFFDu)) {                      // in cartridge space with both operand bytes inside it
            uint32_t ops;
            m.fetch((bank << 12) | (a13 & gp.rom_mask), w, ops);
            b1 = ops & 0xFF; b2 = (ops >> 8) & 0xFF;            // (bits 16..31: case number of a translated core)
        } else {
            w = m.opinfo(peek(pc)); b1 = peek(pc + 1); b2 = peek(pc + 2);
        }
        return exec_body<false>(w, b1, b2);
    }
    // Everything after the fetch.  kConst = the caller passes compile-time constants (a translated
    // cartridge, csrc/gen/xl_*.cuh): decode, addressing and dispatch then fold away at compile time and
    // what is left is the few instructions this one 6502 instruction needs.
    // kBlock (translated cores only) = the instruction is part of a straight-line block: it retires without
    // the end-of-line test (the translator places that test at block ends, and sends an instruction that may
    // touch TIA / RIOT beyond the end of the line to the interpreter beforehand), and the return value is
    // "branch taken" for a branch, false otherwise.  Running CPU-and-RAM-only instructions a little past the
    // end of a scanline before the line is wrapped changes nothing: end_line() reads no CPU or RAM state.
    template <bool kConst, bool kBlock = false>
    STM_DEV bool exec_body(const uint32_t w, uint32_t b1, uint32_t b2)
    {
        const uint32_t am = w & 7, kind = (w >> 5) & 3, grp = (w >> 13) & 31, data = w >> 22;
        uint32_t cyc = (w >> 7) & 15;
        uint32_t npc = pc + ((w >> 11) & 3);                 // pc may run past 0xFFFF: every use masks it

        // effective address
        uint32_t ea;
        {
            uint32_t idx = (w & 8) ? x : ((w & 16) ? y : 0u);
            if (STM_LIKELY(am <= 3)) {                         // none / imm / zp-family / abs-family: branch-free
                uint32_t base = b1 | (am == 3 ? b2 << 8 : 0u);
                uint32_t full = base + idx;
                ea = full & (am == 3 ? 0xFFFFu : 0xFFu);
                cyc += (w >> 18) & ((full ^ base) >> 8) & 1u;
            } else if (am == 4) { ea = peek16zp(b1 + x); }
            else if (am == 5) {
                uint32_t base = peek16zp(b1);
                ea = (base + y) & 0xFFFF;
                if (kind == K_R && ((ea ^ base) & 0xFF00)) cyc++;
            } else if (am == 6) {
                uint32_t ptr = b1 | (b2 << 8);
                ea = peek(ptr) | (peek((ptr & 0xFF00) | ((ptr + 1) & 0xFF)) << 8);
            } else ea = 0;
        }
        if (grp == G_BRANCH) {
            uint32_t fsel = data & 3;
            bool fset = fsel == 0 ? (zn & 0x180u) != 0 : (fsel == 1 ? (p & FV) != 0 : (fsel == 2 ? (p & FC) != 0 : (zn & 0xFFu) == 0));
            const bool taken = fset == ((data & 4) != 0);
            if (taken) {
                uint32_t t = (npc + (uint32_t)(int32_t)(int8_t)(uint8_t)b1) & 0xFFFF;
                cyc += 1 + (((t ^ npc) & 0xFF00) ? 1u : 0u);
                npc = t;
            }
            pc = npc;                                          // nothing else to do for a branch: retire
            hcyc += (int32_t)cyc;
            timer -= (int32_t)cyc;
            if (kBlock) return taken;
            return hcyc >= STM_CYC_PER_LINE;                   // (a branch cannot end the frame)
        }

        // A translated instruction (kConst) knows its base address: an indexed access whose whole 256-byte
        // range lies in cartridge space (no hotspot) is a ROM read, a zero-page one is RAM-or-TIA; the other
        // address spaces' code is then never instantiated for it.  Stack pulls / pushes go the general way.
        uint32_t acls = A_ANY;
        if (kConst && !((w >> 19) & 1u) && grp != G_PUSH && grp != G_JSR && grp != G_BRK) {
            if (am == 2) acls = A_ZP;
            else if (am == 3 && (w & 0x18)) {
                const uint32_t base16 = b1 | (b2 << 8), lo = base16 & 0x1FFF, hi = lo + 0xFF;
                const bool f8 = gp.bank_scheme == STM_BANK_F8;
                if ((lo & 0x1000) && hi <= 0x1FFF && !(f8 && hi >= 0x1FF8 && lo <= 0x1FF9) && kind == K_R) acls = A_ROM;
            }
        }

        // read phase: one memory operand, or 1..3 stack pulls
        uint32_t r0 = b1, r1 = 0, r2 = 0;
        if (w & (3u << 20)) {
            uint32_t nrd = (w >> 20) & 3u;
            uint32_t addr = ea;
            if (STM_UNLIKELY(w & (1u << 19))) { sp = (sp + 1) & 0xFF; addr = 0x100 | sp; }
            for (;;) {                                         // runs once except for RTS / RTI
                r0 = bus_read<kConst>(addr, cyc, w, acls);
                if (STM_LIKELY(--nrd == 0)) break;
                r2 = r1; r1 = r0;
                sp = (sp + 1) & 0xFF; addr = 0x100 | sp;
            }
        }
        const uint32_t v = r0;

        // execute: registers only; memory results are queued in w0..w2
        uint32_t nwr = 0, w0 = 0, w1 = 0, w2 = 0;
        bool wpush = false;
        // The commonest groups are tested one by one (two instructions each, on a copy of the group
        // number the compiler cannot see through — otherwise it folds the chain back into the jump
        // table, whose indexed constant load + indirect branch cost more than the whole chain); the
        // rest go through the jump table.
        auto is = [&](uint32_t g) { uint32_t t = grp; if (!kConst) STM_OPAQUE(t); return t == g; };   // a fresh copy per test
        if (is(G_STORE)) {
            w0 = ((data & 1) ? a : 0xFFu) & ((data & 2) ? x : 0xFFu) & ((data & 4) ? y : 0xFFu);
            nwr = 1;
        } else if (is(G_LOAD)) {
            if (data & 1) a = v;
            if (data & 2) x = v;
            if (data & 4) y = v;
            zn = v;
        } else if (is(G_CMP)) {
            cmp(data == 0 ? a : (data == 1 ? x : y), v);
        } else if (is(G_INCDEC_REG)) {
            uint32_t d = data >> 2;
            if (data & 1) { x = (x + d) & 0xFF; zn = x; } else { y = (y + d) & 0xFF; zn = y; }
        } else if (is(G_XFER)) {
            uint32_t s = data & 3, d = (data >> 2) & 3;
            uint32_t val = s == 0 ? a : (s == 1 ? x : (s == 2 ? y : sp));
            if (d == 0) a = val; else if (d == 1) x = val; else if (d == 2) y = val; else sp = val;
            if (data & 16) zn = val;
        } else if (is(G_FLAG)) {
            p = (data & 256) ? (p | (data & 0xFF)) : (p & ~(data & 0xFF));
        } else switch (grp) {
        case G_PUSH: w0 = data ? (get_p() | FB) : a; nwr = 1; wpush = true; break;
        case G_PULL: if (data) set_p(r0); else { a = r0; zn = r0; } break;
        case G_LOGIC: a = data == 0 ? (a & v) : (data == 1 ? (a | v) : (a ^ v)); zn = a; break;
        case G_ADC: adc(v); break;
        case G_SBC: sbc(v); break;
        case G_BIT: p = (p & ~(uint32_t)FV) | (v & FV); zn = (v & a) | ((v & 0x80) << 1); break;
        case G_SHIFT: {
            uint32_t r = shift(am == 0 ? a : v, data);
            if (am == 0) a = r; else { w0 = r; nwr = 1; }
            break; }
        case G_INCDEC_MEM: w0 = (v + data) & 0xFF; zn = w0; nwr = 1; break;
        case G_JMP: npc = ea; break;
        case G_JSR: { uint32_t ret = (pc + 2) & 0xFFFF; w0 = ret >> 8; w1 = ret & 0xFF; nwr = 2; wpush = true; npc = ea; break; }
        case G_RTS: npc = (((r0 << 8) | r1) + 1) & 0xFFFF; break;          // r1 = low byte (pulled first), r0 = high
        case G_RTI: set_p(r2); npc = (r0 << 8) | r1; break;
        case G_BRK: {
            uint32_t ret = (pc + 2) & 0xFFFF;
            w0 = ret >> 8; w1 = ret & 0xFF; w2 = get_p() | FB; nwr = 3; wpush = true;
            p |= FI;
            npc = peek(0xFFFE) | (peek(0xFFFF) << 8);
            break; }
        case G_RMW_ALU:
            switch (data) {
            case 0: w0 = shift(v, 0); a |= w0; zn = a; break;                               // SLO
            case 1: w0 = shift(v, 2); a &= w0; zn = a; break;                               // RLA
            case 2: w0 = shift(v, 1); a ^= w0; zn = a; break;                               // SRE
            case 3: w0 = shift(v, 3); adc(w0); break;                                       // RRA
            case 4: w0 = (v - 1) & 0xFF; cmp(a, w0); break;                                 // DCP
            default: w0 = (v + 1) & 0xFF; sbc(w0); break;                                   // ISC
            }
            nwr = 1; break;
        case G_ANC: a &= v; zn = a; p = (p & ~(uint32_t)FC) | (a >> 7); break;
        case G_ALR: a = shift(a & v, 1); break;
        case G_SBX: { uint32_t t = (a & x) - v; p = (p & ~(uint32_t)FC) | (((t >> 8) & 1u) ^ 1u); x = t & 0xFF; zn = x; break; }
        default: break;     // G_NOP; the groups resolved above
        }

        // write phase
        if (nwr) {
            uint32_t addr = ea;
            if (STM_UNLIKELY(wpush)) { addr = 0x100 | sp; sp = (sp - 1) & 0xFF; }
            for (;;) {                                         // runs once except for JSR / BRK
                bus_write<kConst>(addr, w0, cyc, acls);
                if (STM_LIKELY(--nwr == 0)) break;
                w0 = w1; w1 = w2;
                addr = 0x100 | sp; sp = (sp - 1) & 0xFF;
            }
        }

        // retire
        pc = npc;
        hcyc += (int32_t)cyc;
        timer -= (int32_t)cyc;
        if (kBlock) return false;
        return hcyc + brk >= STM_CYC_PER_LINE;
    }
    STM_DEV void finish_line()
    {
        if (hcyc >= STM_CYC_PER_LINE) { end_line(); hcyc -= STM_CYC_PER_LINE; }
        if ((brk & BRK_WSYNC) && hcyc != 0) { timer -= STM_CYC_PER_LINE - hcyc; end_line(); hcyc = 0; }
        brk &= ~BRK_WSYNC;
    }
    STM_DEV void exec_one() { exec_core(); if (hcyc >= STM_CYC_PER_LINE || (brk & BRK_WSYNC)) finish_line(); }

    // ---------------- one emulated frame ----------------
    // Rows the frame never reached are zero-filled, as the oracle does.
    // Scanline lock-step: each lane runs instructions until its scanline ends, then the warp — on the GPU the warps of the
    // block that share a scheduler, `lanes` = their threads and barrier (stm_group_all) — reconverges and wraps the line (the whole-line catch-up
    // render) with all lanes active; lanes whose frame is over idle until the slowest lane's frame is over.  Every thread of
    // the group calls run_frame the same number of times, dead lanes (`live` false) included.
    template <class Core = InterpCore>
    STM_DEV void run_frame(uint32_t ale_action, bool want_pixels, uint32_t lanes = 32u, bool live = true)
    {
        (void)lanes;
        if (live) {
            render = want_pixels;
            set_action(ale_action);
            flags &= ~(uint32_t)STM_F_FRAME_DONE;
            brk &= ~BRK_DONE;
            tia_dirty();                                       // a new native frame: no row to reuse yet
        }
        bool over = !live;
        do {
            // Run every lane to the end of its scanline: Core::run_line (the interpreter loop below, or a
            // translated cartridge).
            Core::run_line(*this, lanes, over);
            bool ended;
            ended = hcyc >= STM_CYC_PER_LINE || (brk & BRK_WSYNC);
            if (!over) {
                if (ended) finish_line();
                over = (flags & STM_F_FRAME_DONE) != 0;
            }
        } while (!STM_LINE_ALL(lanes, over));
        if (!live) return;
        if (want_pixels) {
            int r0 = (int)end_line_no - STM_YSTART; if (r0 < 0) r0 = 0;
            for (int r = r0; r < STM_VIS_H; r++)
                if (row_sampled(r)) m.zero_row((uint32_t)r);
        }
        m.st32(O_FRAME, m.ld32(O_FRAME) + 1);
    }

    // ---------------- get_data (cule_bfs.py:157-158) ----------------
    STM_DEV int bcd_score(const uint8_t *addr)
    {
        int s = 0;
        for (int i = 0; i < 3; i++) {
            if (addr[i] == 0xFF) continue;
            uint32_t b = m.ld8(addr[i] & 0x7F);
            s = s * 100 + (int)(b >> 4) * 10 + (int)(b & 15);
        }
        return s;
    }
    STM_DEV int game_lives()
    {
        if (gp.lives_addr == 0xFF) return 0;
        return (int)((m.ld8(gp.lives_addr & 0x7F) >> gp.lives_shift) & gp.lives_mask);
    }
    // returns the frame reward; updates done / lives
    STM_DEV int get_data(bool episodic_life, uint32_t &done, int &lives)
    {
        int score = bcd_score(gp.score_addr) - bcd_score(gp.score2_addr);
        int r = score - (int)m.ld32(O_SCORE_PREV);
        m.st32(O_SCORE_PREV, (uint32_t)score);
        int lv = game_lives();
        bool term = false;
        if (gp.term_addr != 0xFF) term = (m.ld8(gp.term_addr & 0x7F) & gp.term_mask) == gp.term_value;
        if (term || (episodic_life && lv < lives)) done = 1;
        lives = lv;
        return r;
    }
};

} // namespace stm
