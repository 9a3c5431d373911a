// host_core.cpp — TEST AID.  Compiles softtreemax_b200/csrc/emu_core.cuh for the host with a
// plain-array memory policy so that the CUDA emulator core can be diffed against the oracle in
// this GPU-less container (tests/test_host_core.py).  Never loaded by the product package.
#include <cstring>
#include <vector>
#include <cstdint>
#include "../../include/stm_palette.h"
#include "../../softtreemax_b200/csrc/emu_core.cuh"
#include "../../softtreemax_b200/csrc/gen/xl_index.cuh"

namespace {
struct HostMem {
    uint8_t *blob; const uint8_t *rom; const uint32_t *lut; uint8_t *fb;
    uint32_t ld8(uint32_t o) const { return blob[o]; }
    void st8(uint32_t o, uint32_t v) { blob[o] = (uint8_t)v; }
    uint32_t ld16(uint32_t o) const { return blob[o] | (blob[o + 1] << 8); }
    void st16(uint32_t o, uint32_t v) { blob[o] = (uint8_t)v; blob[o + 1] = (uint8_t)(v >> 8); }
    uint32_t ld32(uint32_t o) const { uint32_t v; memcpy(&v, blob + o, 4); return v; }
    void st32(uint32_t o, uint32_t v) { memcpy(blob + o, &v, 4); }
    uint32_t rom8(uint32_t i) const { return rom[i]; }
    uint32_t mask;
    const uint16_t *xl = nullptr;          // case number per image offset (translated cores), or null
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
    void copy_row(uint32_t dst, uint32_t src) { if (fb) memcpy(fb + dst * 160, fb + src * 160, 160); }
    void open_row(uint32_t) {}
    void link_row(uint32_t dst, uint32_t src) { copy_row(dst, src); }
    void zero_row(uint32_t row) { if (fb) memset(fb + row * 160, 0, 160); }
    void px1(uint32_t row, uint32_t x, uint32_t v) { if (fb) fb[row * 160 + x] = (uint8_t)v; }
};
uint32_t g_lut[256];
bool g_lut_ok = false;
}

extern "C" {

void hc_fill_params(stm::GameParams *gp, const stm_game_desc *d)
{
    memset(gp, 0, sizeof(*gp));
    gp->rom_mask = d->bank_scheme == STM_BANK_F8 ? 0xFFF : d->rom_size - 1;
    gp->bank_scheme = d->bank_scheme; gp->controller = d->controller; gp->n_actions = d->n_actions;
    memcpy(gp->actions, d->actions, 18);
    memcpy(gp->score_addr, d->score_addr, 3); memcpy(gp->score2_addr, d->score2_addr, 3);
    gp->lives_addr = d->lives_addr; gp->lives_shift = d->lives_shift; gp->lives_mask = d->lives_mask;
    gp->term_addr = d->term_addr; gp->term_mask = d->term_mask; gp->term_value = d->term_value;
}

void hc_decode_table(uint32_t *out) { stm::build_decode_table(out); }

}  // extern "C"

extern "C" void hc_fill_params(stm::GameParams *gp, const stm_game_desc *d);

// per translated core: image offset -> case number (what stm_ctx_create stamps into the predecoded cartridge)
template <class Core> static const uint16_t *case_table(uint32_t rom_size)
{
    static std::vector<uint16_t> t;
    if (t.empty()) {
        t.assign(rom_size, 0);
        for (uint32_t k = 0; k < Core::kCases; k++) t[Core::offsets()[k]] = (uint16_t)(k + 1);
    }
    return t.data();
}
template <> const uint16_t *case_table<stm::InterpCore>(uint32_t) { return nullptr; }

// one emulated frame + get_data, like the oracle's orc_frame + orc_get_data
template <class Core>
static void frame_with(const stm_game_desc *d, const uint8_t *rom, uint8_t *blob, int ale_action, uint8_t *fb,
                       int episodic_life, float weight, uint8_t *done, float *reward, int32_t *lives)
{
    if (!g_lut_ok) { stm::build_hot_table(g_lut); g_lut_ok = true; }
    stm::GameParams gp; hc_fill_params(&gp, d);
    HostMem m{blob, rom, g_lut, fb, gp.rom_mask};
    m.xl = case_table<Core>(d->rom_size);
    stm::Emu<HostMem> e(m, gp);
    e.load();
    e.template run_frame<Core>((uint32_t)ale_action, fb != nullptr);
    uint32_t dn = *done; int lv = *lives;
    int r = e.get_data(episodic_life != 0, dn, lv);
    volatile float wr = weight * (float)r;
    *reward = *reward + wr;
    *done = (uint8_t)dn; *lives = lv;
    e.store();
}

extern "C" {

void hc_frame(const stm_game_desc *d, const uint8_t *rom, uint8_t *blob, int ale_action, uint8_t *fb,
              int episodic_life, float weight, uint8_t *done, float *reward, int32_t *lives)
{ frame_with<stm::InterpCore>(d, rom, blob, ale_action, fb, episodic_life, weight, done, reward, lives); }

// the same through the translated core of this cartridge (csrc/gen/xl_*.cuh); returns 0 if there is none
int hc_frame_xl(const stm_game_desc *d, const uint8_t *rom, uint8_t *blob, int ale_action, uint8_t *fb,
                int episodic_life, float weight, uint8_t *done, float *reward, int32_t *lives)
{
    const uint32_t h = stm::rom_hash(rom, d->rom_size);
#define STM_TRY(name) \
    if (h == stm::Xl_##name::kRomHash && d->rom_size == stm::Xl_##name::kRomSize) { \
        frame_with<stm::Xl_##name>(d, rom, blob, ale_action, fb, episodic_life, weight, done, reward, lives); return 1; }
    STM_XL_LIST(STM_TRY)
#undef STM_TRY
    return 0;
}

void hc_exec_one(const stm_game_desc *d, const uint8_t *rom, uint8_t *blob)
{
    if (!g_lut_ok) { stm::build_hot_table(g_lut); g_lut_ok = true; }
    stm::GameParams gp; hc_fill_params(&gp, d);
    HostMem m{blob, rom, g_lut, nullptr, gp.rom_mask};
    stm::Emu<HostMem> e(m, gp);
    e.load(); e.exec_one(); e.store();
}

}
