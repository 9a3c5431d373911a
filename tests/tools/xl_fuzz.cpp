// xl_fuzz.cpp — TEST AID (tests/test_xlate_fuzz.py).  Compiles the device emulator core for the host together with ONE
// freshly translated cartridge (tools/xlate_rom.cpp -> XL_HEADER, on the include path) and runs interpreter and
// translation side by side from caller-supplied states: both must leave the same 256-byte state after every frame.
// Never loaded by the product package.
#include <cstring>
#include <cstdint>
#include <vector>
#include "../../include/stm_palette.h"
#include "../../softtreemax_b200/csrc/emu_core.cuh"
#include XL_HEADER        /* -DXL_HEADER=\"xl_<name>.cuh\" -DXL_CORE=Xl_<name>: a name of its own per cartridge — libraries loaded into one
                            process share the static tables of same-named inline functions */

namespace {
struct HostMem {
    uint8_t *blob; const uint8_t *rom; const uint32_t *lut; uint8_t *fb;
    uint32_t mask;
    const uint16_t *xl = nullptr;
    uint32_t ld8(uint32_t o) const { return blob[o]; }
    void st8(uint32_t o, uint32_t v) { blob[o] = (uint8_t)v; }
    uint32_t ld16(uint32_t o) const { return blob[o] | (blob[o + 1] << 8); }
    void st16(uint32_t o, uint32_t v) { blob[o] = (uint8_t)v; blob[o + 1] = (uint8_t)(v >> 8); }
    uint32_t ld32(uint32_t o) const { uint32_t v; memcpy(&v, blob + o, 4); return v; }
    void st32(uint32_t o, uint32_t v) { memcpy(blob + o, &v, 4); }
    uint32_t rom8(uint32_t i) const { return rom[i]; }
    void fetch(uint32_t off, uint32_t &w, uint32_t &ops) const
    {
        uint32_t win = off & ~mask;
        w = lut[rom[off]];
        ops = rom[win | ((off + 1) & mask)] | (rom[win | ((off + 2) & mask)] << 8) | (xl ? (uint32_t)xl[off] << 16 : 0u);
    }
    uint32_t opinfo(uint32_t op) const { return lut[op]; }
    uint32_t grey(uint32_t i) const { static const uint8_t g[128] = STM_GRAY_LUT_INIT; return g[i & 127]; }
    void px4(uint32_t row, uint32_t g, uint32_t w) { if (fb) memcpy(fb + row * 160 + 4 * g, &w, 4); }
    void px16(uint32_t row, uint32_t q, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
    { px4(row, 4 * q, a); px4(row, 4 * q + 1, b); px4(row, 4 * q + 2, c); px4(row, 4 * q + 3, d); }
    void open_row(uint32_t) {}
    void link_row(uint32_t dst, uint32_t src) { if (fb) memcpy(fb + dst * 160, fb + src * 160, 160); }
    void zero_row(uint32_t row) { if (fb) memset(fb + row * 160, 0, 160); }
    void px1(uint32_t row, uint32_t x, uint32_t v) { if (fb) fb[row * 160 + x] = (uint8_t)v; }
};
}  // namespace

extern "C" {

// states: n x 256 (modified in place by the interpreter run).  Returns -1 when every env agrees after every frame, else
// env * 1000 + frame of the first disagreement (state blob or drawn frame).
int xf_run(const stm_game_desc *d, const uint8_t *rom, uint8_t *states, int n, int nframes, const uint8_t *actions)
{
    static uint32_t lut[256]; static bool ok = false;
    if (!ok) { stm::build_hot_table(lut); ok = true; }
    stm::GameParams gp; memset(&gp, 0, sizeof gp);
    gp.rom_mask = d->bank_scheme == STM_BANK_F8 ? 0xFFF : d->rom_size - 1;
    gp.bank_scheme = d->bank_scheme; gp.controller = d->controller; gp.n_actions = d->n_actions;
    memcpy(gp.actions, d->actions, 18);
    memcpy(gp.score_addr, d->score_addr, 3); memcpy(gp.score2_addr, d->score2_addr, 3);
    gp.lives_addr = d->lives_addr; gp.lives_shift = d->lives_shift; gp.lives_mask = d->lives_mask;
    gp.term_addr = d->term_addr; gp.term_mask = d->term_mask; gp.term_value = d->term_value;
    if (stm::rom_hash(rom, d->rom_size) != stm::XL_CORE::kRomHash) return -2;
    std::vector<uint16_t> ct(d->rom_size, 0);
    for (uint32_t k = 0; k < stm::XL_CORE::kCases; k++) ct[stm::XL_CORE::offsets()[k]] = (uint16_t)(k + 1);
    std::vector<uint8_t> fa(33600), fb(33600);
    for (int env = 0; env < n; env++) {
        uint8_t *a = states + (size_t)env * 256, b[256];
        memcpy(b, a, 256);
        for (int f = 0; f < nframes; f++) {
            const bool draw = (f & 1) != 0;
            HostMem ma{a, rom, lut, draw ? fa.data() : nullptr, gp.rom_mask}, mb{b, rom, lut, draw ? fb.data() : nullptr, gp.rom_mask};
            mb.xl = ct.data();
            stm::Emu<HostMem> ea(ma, gp), eb(mb, gp);
            ea.load(); eb.load();
            ea.run_frame<stm::InterpCore>(actions[env], draw);
            eb.run_frame<stm::XL_CORE>(actions[env], draw);
            ea.store(); eb.store();
            if (memcmp(a, b, 256) != 0 || (draw && fa != fb)) return env * 1000 + f;
        }
    }
    return -1;
}

}
