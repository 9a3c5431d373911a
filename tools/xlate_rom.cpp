// xlate_rom — ahead-of-time translator of a cartridge for the K2 step kernel.
//
//   xlate_rom <rom.bin> <name> <bank_scheme: 0 none | 1 F8> <out.cuh>
//
// Walks the cartridge's code from its vectors (recursive descent over branches, JMP/JSR targets, return
// sites, both banks across an F8 hotspot access) and emits a header in which every reachable instruction is
// a call of Emu::exec_body<true, true> with the instruction's hot decode word and operand bytes as
// COMPILE-TIME constants, so the compiler folds decode, addressing and dispatch away and keeps the few
// instructions that 6502 instruction needs.  The semantics are exec_body's, i.e. the interpreter's — the
// translated core is bit-identical to it by construction; anything the walk did not reach (computed jumps,
// code in RAM, operands straddling the bank window) goes to the `default:` case = the interpreter.
//
// Block structure.  Consecutive instructions form straight-line BLOCKS: inside a block nothing is tested
// between instructions (so the compiler also merges their cycle / program-counter updates and drops flag
// results nobody reads); the end-of-scanline test sits at block ends only — branches, jumps, calls, returns,
// a strobe of WSYNC / VSYNC, and every 60 emulated cycles.  That is exact: an instruction that touches only
// CPU registers, RAM or ROM has the same effect a few cycles before or after the scanline is wrapped, and an
// instruction that may touch TIA or RIOT is first checked against the end of the line (`hcyc > 76 - cycles`)
// and handed to the interpreter, which wraps the line at the exact cycle, when it would reach it.
// Control flow between blocks is direct (`goto` to a static target); only block entry points are `case`
// labels of the dispatch switch (offsets() lists them), reached after a scanline wrap, a return, an
// indirect jump or an interpreted instruction.  In the last lines before the frame-length cap everything is
// interpreted, so that a frame cut off by the cap ends at the exact instruction.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iterator>
#include <set>
#include <vector>
#include "../softtreemax_b200/csrc/emu_core.cuh"

// does [a16, a16 + reach] touch an F8 bank-switch hotspot ($1FF8 / $1FF9 and their mirrors)?
static bool spans_hotspot(uint32_t a16, uint32_t reach)
{
    for (uint32_t k = 0; k <= reach; k++) { const uint32_t t = ((a16 + k) & 0xFFFFu) & 0x1FFFu; if (t == 0x1FF8u || t == 0x1FF9u) return true; }
    return false;
}

int main(int argc, char **argv)
{
    if (argc != 5 && argc != 6) { fprintf(stderr, "usage: xlate_rom <rom.bin> <name> <bank_scheme> <out.cuh> [max instructions]\n"); return 2; }
    // an optional cap on the instructions translated (compile time grows faster than linearly with them); what the walk
    // does not reach before it is interpreted — every jump into untranslated code goes through the dispatcher anyway
    const size_t max_ins = argc == 6 ? (size_t)atoi(argv[5]) : 6000u;
    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 1; }
    std::vector<uint8_t> rom(65536);
    size_t n = fread(rom.data(), 1, rom.size(), f);
    fclose(f);
    rom.resize(n);
    const char *name = argv[2];
    const int f8 = atoi(argv[3]);
    if (!((f8 == 0 && (n == 2048 || n == 4096)) || (f8 == 1 && n == 8192))) { fprintf(stderr, "unsupported size %zu\n", n); return 1; }
    const uint32_t mask = f8 ? 0xFFFu : (uint32_t)n - 1, nbanks = f8 ? 2 : 1;
    uint32_t hot[256];
    stm::build_hot_table(hot);

    uint32_t hash = 2166136261u;                                   // FNV-1a over the image
    for (uint8_t b : rom) { hash ^= b; hash *= 16777619u; }

    auto at = [&](uint32_t bank, uint32_t a13) -> uint32_t { return (bank << 12) | (a13 & mask); };   // image offset of a bus address
    std::set<uint32_t> seen, labels;
    std::vector<uint32_t> work;
    auto push_addr = [&](uint32_t bank, uint32_t addr16, bool label) {        // only cartridge space
        if (!(addr16 & 0x1000)) return;
        uint32_t off = at(bank, addr16 & 0x1FFF);
        if (label) labels.insert(off);
        if (seen.insert(off).second) work.push_back(off);
    };
    for (uint32_t b = 0; b < nbanks; b++)
        for (uint32_t vec : { 0xFFFCu, 0xFFFEu })
            push_addr(b, rom[at(b, vec)] | (rom[at(b, vec + 1)] << 8), true);
    std::vector<uint32_t> order;
    while (!work.empty()) {
        uint32_t off = work.back(); work.pop_back();
        uint32_t bank = off >> 12, win = off & mask;
        if (win > mask - 2) { seen.erase(off); continue; }       // operand bytes would wrap the window: interpreter
        order.push_back(off);
        uint32_t op = rom[off], w = hot[op], b1 = rom[off + 1], b2 = rom[off + 2];
        uint32_t len = (w >> 11) & 3, grp = (w >> 13) & 31, am = w & 7;
        uint32_t pc = 0x1000 | win;                                // a representative bus address of this offset
        uint32_t next = pc + len;
        auto both_banks = [&](uint32_t addr16) { for (uint32_t b = 0; b < nbanks; b++) push_addr(b, addr16, true); };
        switch (grp) {
        case stm::G_BRANCH: push_addr(bank, next, true); push_addr(bank, (next + (uint32_t)(int32_t)(int8_t)b1) & 0x1FFF, true); break;
        case stm::G_JMP:
            if (am == 3) push_addr(bank, b1 | (b2 << 8), true);
            else { uint32_t p = b1 | (b2 << 8); if (p & 0x1000) push_addr(bank, rom[at(bank, p)] | (rom[at(bank, (p & 0xFF00) | ((p + 1) & 0xFF))] << 8), true); }
            break;
        case stm::G_JSR: push_addr(bank, b1 | (b2 << 8), true); push_addr(bank, next, true); break;
        case stm::G_RTS: case stm::G_RTI: break;
        case stm::G_BRK: push_addr(bank, rom[at(bank, 0xFFFE)] | (rom[at(bank, 0xFFFF)] << 8), true); push_addr(bank, pc + 2, true); break;
        default: {
            // an absolute operand on an F8 hotspot: the next instruction runs in the other bank — walk both.  (Indexed and
            // indirect operands may land there too; those are caught at run time — see dynswitch — and what they reach in the
            // other bank is interpreted unless something else leads the walk there.)
            const uint32_t idx = (w >> 3) & 3;
            bool hotspot = f8 && am == 3 && !idx && spans_hotspot(b1 | (b2 << 8), 0u);
            if (hotspot) both_banks(next); else push_addr(bank, next, false);
            break; }
        }
        if (order.size() >= max_ins) { for (uint32_t o : work) seen.erase(o); work.clear(); }
    }
    std::set<uint32_t> cases(order.begin(), order.end());

    // per-instruction facts
    struct Ins { uint32_t off, w, b1, b2, len, grp, am, idx, kind, cyc, cmax, next; bool touch, strobe, hotspot, dynswitch, next_ok; };
    std::vector<Ins> ins;
    for (auto it = cases.begin(); it != cases.end(); ++it) {
        Ins q{};
        q.off = *it; q.w = hot[rom[q.off]]; q.b1 = rom[q.off + 1]; q.b2 = rom[q.off + 2];
        q.am = q.w & 7; q.idx = (q.w >> 3) & 3; q.kind = (q.w >> 5) & 3; q.cyc = (q.w >> 7) & 15; q.len = (q.w >> 11) & 3; q.grp = (q.w >> 13) & 31;
        q.next = (q.off & ~mask) | ((q.off + q.len) & mask);
        auto nx = std::next(it);
        q.next_ok = nx != cases.end() && *nx == q.next && ((q.off & mask) + q.len <= mask);
        q.cmax = q.cyc + (((q.w >> 18) & 1) ? 1 : 0) + ((q.am == 5 && q.kind == stm::K_R) ? 1 : 0) + (q.grp == stm::G_BRANCH ? 2 : 0);
        const uint32_t a16 = q.b1 | (q.b2 << 8), a13 = a16 & 0x1FFF;
        // may the instruction touch TIA / RIOT?  (=> it must not straddle the end of the line inside a block)
        const bool stack = q.grp == stm::G_PUSH || q.grp == stm::G_PULL || q.grp == stm::G_JSR || q.grp == stm::G_RTS || q.grp == stm::G_RTI || q.grp == stm::G_BRK;
        bool touch = false, tia_static = false; uint32_t tia_reg = 0;
        if (stack) touch = true;
        else if ((q.kind & 3) && q.am != 1 && q.am != 0 && q.am != 7) {
            if (q.am == 2 && !q.idx) { touch = q.b1 < 0x80; tia_static = touch; tia_reg = q.b1 & 0x3F; }
            else if (q.am == 3 && !q.idx) {
                touch = !(a13 & 0x1000) && (a13 & 0x1280) != 0x80;
                if (touch && !(a13 & 0x80)) { tia_static = true; tia_reg = a13 & 0x3F; }
            }
            else if (q.am == 3 && q.kind == stm::K_R) { uint32_t lo = a13, hi = lo + 0xFF;
                                                      touch = !((lo & 0x1000) && hi <= 0x1FFF && !(f8 && hi >= 0x1FF8 && lo <= 0x1FF9)); }
            else touch = true;
        }
        q.touch = touch;
        // may it strobe WSYNC / VSYNC (the only writes that raise a break request)?
        const bool writes = (q.kind & 2) != 0 || q.grp == stm::G_PUSH || q.grp == stm::G_JSR || q.grp == stm::G_BRK;
        q.strobe = touch && writes && (!tia_static || tia_reg == 0x00 || tia_reg == 0x02);
        if (touch && writes && !tia_static && !stack && q.am == 3 && !q.idx && (a13 & 0x80)) q.strobe = false;   // static RIOT write
        q.hotspot = f8 && q.am == 3 && !q.idx && (a13 == 0x1FF8 || a13 == 0x1FF9);
        // operands only known at run time that may land on a hotspot (indexed absolute reaching it, any indirect one): the block
        // goes on, but the bank is compared with the one it was translated for right after the instruction
        q.dynswitch = f8 && !q.hotspot && ((q.am == 3 && q.idx && spans_hotspot(a16, 255u)) || q.am == 4 || q.am == 5);
        ins.push_back(q);
    }
    // entry points: control-flow targets (collected above), the instruction after a possible strobe, starts of runs,
    // and a split wherever a straight run exceeds 60 emulated cycles without an end-of-line test
    {
        uint32_t run = 0;
        for (size_t i = 0; i < ins.size(); i++) {
            const Ins &q = ins[i];
            const bool first = i == 0 || !ins[i - 1].next_ok || ins[i - 1].next != q.off;
            if (first) labels.insert(q.off);
            if (i > 0 && ins[i - 1].next_ok && ins[i - 1].strobe) labels.insert(q.off);
            if (labels.count(q.off)) run = 0;
            if (run + q.cmax > 60) { labels.insert(q.off); run = 0; }
            run += q.cmax;
        }
    }
    std::vector<uint32_t> label_list;
    for (uint32_t off : labels) if (cases.count(off)) label_list.push_back(off);
    std::set<uint32_t> label_set(label_list.begin(), label_list.end());
    if (label_list.size() > 65535) { fprintf(stderr, "too many entry points\n"); return 1; }

    FILE *o = fopen(argv[4], "w");
    if (!o) { perror(argv[4]); return 1; }
    fprintf(o, "// GENERATED by tools/xlate_rom.cpp from the cartridge image \"%s\" (%zu bytes) — do not edit.\n", name, n);
    fprintf(o, "// %zu reachable instructions translated in straight-line blocks, %zu block entry points; everything else\n", cases.size(), label_list.size());
    fprintf(o, "// runs through the interpreter (default case).\n");
    fprintf(o, "#pragma once\nnamespace stm {\nstruct Xl_%s {\n", name);
    fprintf(o, "    static constexpr uint32_t kRomHash = 0x%08Xu, kRomSize = %zuu, kCases = %zuu;\n", hash, n, label_list.size());
    fprintf(o, "    // image offsets of the block entry points, ascending; case k+1 is kOffsets[k] (the predecoded\n");
    fprintf(o, "    // cartridge carries k+1 in bits 16..31 of the operand word of that offset: stm::stamp_cases)\n");
    fprintf(o, "    static const uint16_t *offsets() {\n        static const uint16_t k[kCases] = {");
    { int i = 0; for (uint32_t off : label_list) fprintf(o, "%s0x%04X,", (i++ % 16) ? " " : "\n            ", off); }
    fprintf(o, "\n        };\n        return k;\n    }\n");
    fprintf(o, "    template <class E>\n    static STM_DEV void run_line(E &e, uint32_t lanes, bool over)\n    {\n");
    fprintf(o, "        (void)lanes;\n        if (over) return;\n        for (;;) {\n");
    fprintf(o, "            uint32_t cs = 0;\n");
    fprintf(o, "            {\n                const uint32_t a13 = e.pc & 0x1FFFu;\n");
    fprintf(o, "                if (a13 - 0x1000u <= 0xFFDu) { uint32_t w, ops; e.m.fetch((e.bank << 12) | (a13 & 0x%Xu), w, ops); cs = ops >> 16; }\n", mask);
    fprintf(o, "            }\n");
    fprintf(o, "            if (e.line >= (uint32_t)STM_MAX_LINES - 3u) cs = 0;            // near the frame-length cap: exact instruction boundaries\n");
    fprintf(o, "            switch (cs) {\n");
    {
        std::vector<uint32_t> case_no(65536, 0);
        for (size_t k = 0; k < label_list.size(); k++) case_no[label_list[k]] = (uint32_t)k + 1;
        // entry points that some block jumps to directly get a goto label as well
        std::set<uint32_t> jumped;
        for (const Ins &q : ins) {
            uint32_t a16;
            if (q.grp == stm::G_BRANCH) a16 = ((0x1000u | (q.off & mask)) + 2u + (uint32_t)(int32_t)(int8_t)q.b1) & 0x1FFFu;
            else if ((q.grp == stm::G_JMP && q.am == 3) || q.grp == stm::G_JSR) a16 = q.b1 | (q.b2 << 8);
            else continue;
            if (a16 & 0x1000) jumped.insert(at(q.off >> 12, a16 & 0x1FFF));
        }
        for (size_t i = 0; i < ins.size(); i++) {
            const Ins &q = ins[i];
            const bool labeled = label_set.count(q.off) != 0;
            const bool flow = q.grp == stm::G_BRANCH;
            const bool leaves = q.grp == stm::G_JMP || q.grp == stm::G_JSR || q.grp == stm::G_RTS || q.grp == stm::G_RTI || q.grp == stm::G_BRK;
            // a run that falls into a labelled successor placed a split there: test the end of the line before crossing it
            if (labeled && jumped.count(q.off)) fprintf(o, "            case %u: L_%04X:                                              // image offset 0x%04X\n", case_no[q.off], q.off, q.off);
            else if (labeled) fprintf(o, "            case %u:                                                      // image offset 0x%04X\n", case_no[q.off], q.off);
            if (q.touch) fprintf(o, "                if (STM_UNLIKELY(e.hcyc > %d)) goto interp;\n", 76 - (int)q.cmax);
            // direct jump to a static target on the bus (same bank) when it is a translated entry point, else the dispatcher
            auto target_goto = [&](uint32_t addr16) {
                const uint32_t toff = at(q.off >> 12, addr16 & 0x1FFF);
                if ((addr16 & 0x1000) && cases.count(toff) && label_set.count(toff)) fprintf(o, "goto L_%04X;", toff); else fprintf(o, "continue;");
            };
            if (flow) {
                const uint32_t t16 = ((0x1000u | (q.off & mask)) + 2u + (uint32_t)(int32_t)(int8_t)q.b1) & 0x1FFFu;   // as the walk above
                fprintf(o, "                if (e.template exec_body<true, true>(0x%08Xu, 0x%02Xu, 0x%02Xu)) { if (e.hcyc >= STM_CYC_PER_LINE) return; ", q.w, q.b1, q.b2);
                target_goto(t16);
                fprintf(o, " }\n");
                fprintf(o, "                if (e.hcyc >= STM_CYC_PER_LINE) return;\n");
                if (!q.next_ok) fprintf(o, "                continue;\n");
                continue;
            }
            fprintf(o, "                e.template exec_body<true, true>(0x%08Xu, 0x%02Xu, 0x%02Xu);\n", q.w, q.b1, q.b2);
            if (q.dynswitch && !leaves)
                fprintf(o, "                if (STM_UNLIKELY(e.bank != %uu)) { if (e.hcyc + e.brk >= STM_CYC_PER_LINE) return; continue; }\n", q.off >> 12);
            if (q.grp == stm::G_JMP && q.am == 3) {
                fprintf(o, "                if (e.hcyc >= STM_CYC_PER_LINE) return;\n                ");
                target_goto(q.b1 | (q.b2 << 8));
                fprintf(o, "\n");
                continue;
            }
            if (q.grp == stm::G_JSR) {
                fprintf(o, "                if (e.hcyc + e.brk >= STM_CYC_PER_LINE) return;\n                ");
                target_goto(q.b1 | (q.b2 << 8));
                fprintf(o, "\n");
                continue;
            }
            if (leaves || q.hotspot || !q.next_ok) {
                fprintf(o, "                if (e.hcyc + e.brk >= STM_CYC_PER_LINE) return;\n                continue;\n");
                continue;
            }
            if (q.strobe) fprintf(o, "                if (e.brk) return;\n");
            // falling into an entry point: the end-of-line test of the block end
            if (label_set.count(q.next)) fprintf(o, "                if (e.hcyc >= STM_CYC_PER_LINE) return;\n");
        }
    }
    fprintf(o, "            default:\n            interp:\n");
    fprintf(o, "                if (e.exec_core()) return;\n");
    fprintf(o, "                break;\n            }\n");
    fprintf(o, "        }\n    }\n};\n}  // namespace stm\n");
    fclose(o);
    fprintf(stderr, "%s: %zu instructions, %zu entry points, hash %08X\n", name, cases.size(), label_list.size(), hash);
    return 0;
}
