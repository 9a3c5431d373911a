"""Tiny two-pass 6502 assembler used at build time to make the synthetic cartridges.

No copyrighted Atari ROM is available offline, so the workloads BASELINE.json
names (Breakout / Pong / MsPacman action-set classes) are stood in for by small games written
here in 6502 assembly.  This file only turns source text into bytes; it is not on any hot path.

Syntax: `label:`, `NAME = expr`, `.org expr`, `.byte e,e,...`, `.word e,...`, `.fill n,v`,
`.align n`, mnemonics with the usual operand forms (`#e`, `e`, `e,X`, `e,Y`, `(e,X)`, `(e),Y`,
`(e)`, `A`).  Expressions are Python with `$hex`, `%bin`, `<e` (low byte), `>e` (high byte), `*`
as a leading token meaning the current address.
"""
import re

_OPS = {
    # mnemonic: {mode: opcode}
    "ADC": {"imm": 0x69, "zp": 0x65, "zpx": 0x75, "abs": 0x6D, "abx": 0x7D, "aby": 0x79, "inx": 0x61, "iny": 0x71},
    "AND": {"imm": 0x29, "zp": 0x25, "zpx": 0x35, "abs": 0x2D, "abx": 0x3D, "aby": 0x39, "inx": 0x21, "iny": 0x31},
    "ASL": {"acc": 0x0A, "zp": 0x06, "zpx": 0x16, "abs": 0x0E, "abx": 0x1E},
    "BIT": {"zp": 0x24, "abs": 0x2C},
    "CMP": {"imm": 0xC9, "zp": 0xC5, "zpx": 0xD5, "abs": 0xCD, "abx": 0xDD, "aby": 0xD9, "inx": 0xC1, "iny": 0xD1},
    "CPX": {"imm": 0xE0, "zp": 0xE4, "abs": 0xEC},
    "CPY": {"imm": 0xC0, "zp": 0xC4, "abs": 0xCC},
    "DEC": {"zp": 0xC6, "zpx": 0xD6, "abs": 0xCE, "abx": 0xDE},
    "EOR": {"imm": 0x49, "zp": 0x45, "zpx": 0x55, "abs": 0x4D, "abx": 0x5D, "aby": 0x59, "inx": 0x41, "iny": 0x51},
    "INC": {"zp": 0xE6, "zpx": 0xF6, "abs": 0xEE, "abx": 0xFE},
    "JMP": {"abs": 0x4C, "ind": 0x6C},
    "JSR": {"abs": 0x20},
    "LDA": {"imm": 0xA9, "zp": 0xA5, "zpx": 0xB5, "abs": 0xAD, "abx": 0xBD, "aby": 0xB9, "inx": 0xA1, "iny": 0xB1},
    "LDX": {"imm": 0xA2, "zp": 0xA6, "zpy": 0xB6, "abs": 0xAE, "aby": 0xBE},
    "LDY": {"imm": 0xA0, "zp": 0xA4, "zpx": 0xB4, "abs": 0xAC, "abx": 0xBC},
    "LSR": {"acc": 0x4A, "zp": 0x46, "zpx": 0x56, "abs": 0x4E, "abx": 0x5E},
    "ORA": {"imm": 0x09, "zp": 0x05, "zpx": 0x15, "abs": 0x0D, "abx": 0x1D, "aby": 0x19, "inx": 0x01, "iny": 0x11},
    "ROL": {"acc": 0x2A, "zp": 0x26, "zpx": 0x36, "abs": 0x2E, "abx": 0x3E},
    "ROR": {"acc": 0x6A, "zp": 0x66, "zpx": 0x76, "abs": 0x6E, "abx": 0x7E},
    "SBC": {"imm": 0xE9, "zp": 0xE5, "zpx": 0xF5, "abs": 0xED, "abx": 0xFD, "aby": 0xF9, "inx": 0xE1, "iny": 0xF1},
    "STA": {"zp": 0x85, "zpx": 0x95, "abs": 0x8D, "abx": 0x9D, "aby": 0x99, "inx": 0x81, "iny": 0x91},
    "STX": {"zp": 0x86, "zpy": 0x96, "abs": 0x8E},
    "STY": {"zp": 0x84, "zpx": 0x94, "abs": 0x8C},
    # undocumented but stable, used by real carts
    "LAX": {"zp": 0xA7, "zpy": 0xB7, "abs": 0xAF, "aby": 0xBF, "inx": 0xA3, "iny": 0xB3},
    "SAX": {"zp": 0x87, "zpy": 0x97, "abs": 0x8F, "inx": 0x83},
    "DCP": {"zp": 0xC7, "zpx": 0xD7, "abs": 0xCF, "abx": 0xDF, "aby": 0xDB, "inx": 0xC3, "iny": 0xD3},
    "ISC": {"zp": 0xE7, "zpx": 0xF7, "abs": 0xEF, "abx": 0xFF, "aby": 0xFB, "inx": 0xE3, "iny": 0xF3},
}
_IMPLIED = {
    "BRK": 0x00, "CLC": 0x18, "CLD": 0xD8, "CLI": 0x58, "CLV": 0xB8, "DEX": 0xCA, "DEY": 0x88,
    "INX": 0xE8, "INY": 0xC8, "NOP": 0xEA, "PHA": 0x48, "PHP": 0x08, "PLA": 0x68, "PLP": 0x28,
    "RTI": 0x40, "RTS": 0x60, "SEC": 0x38, "SED": 0xF8, "SEI": 0x78, "TAX": 0xAA, "TAY": 0xA8,
    "TSX": 0xBA, "TXA": 0x8A, "TXS": 0x9A, "TYA": 0x98,
}
_BRANCH = {"BPL": 0x10, "BMI": 0x30, "BVC": 0x50, "BVS": 0x70, "BCC": 0x90, "BCS": 0xB0, "BNE": 0xD0, "BEQ": 0xF0}
_LEN = {"imm": 2, "zp": 2, "zpx": 2, "zpy": 2, "abs": 3, "abx": 3, "aby": 3, "inx": 2, "iny": 2, "ind": 3, "acc": 1}


class AsmError(Exception):
    pass


def _expr(text, syms, pc, strict):
    t = text.strip()
    lo = hi = False
    if t.startswith("<"):
        lo, t = True, t[1:]
    elif t.startswith(">"):
        hi, t = True, t[1:]
    t = re.sub(r"\$([0-9A-Fa-f]+)", lambda m: str(int(m.group(1), 16)), t)
    t = re.sub(r"%([01]+)", lambda m: str(int(m.group(1), 2)), t)
    t = re.sub(r"(^|[(+\-*/&|^ ,])\*(?=$|[+\-) ])", lambda m: m.group(1) + str(pc), t)
    try:
        v = eval(t, {"__builtins__": {}}, syms)
    except NameError:
        if strict:
            raise AsmError("undefined symbol in %r" % text)
        return None
    v = int(v)
    if lo:
        v &= 0xFF
    if hi:
        v = (v >> 8) & 0xFF
    return v


def _split_args(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch == "(":
            depth += 1
        if ch == ")":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur)
    return out


def assemble(src, origin=0xF000, size=4096, fill=0xFF, predefined=None):
    """Return (bytes, symbols). `origin` is the address of image byte 0."""
    lines = []
    for raw in src.splitlines():
        line = raw.split(";")[0].rstrip()
        if line.strip():
            lines.append(line)
    syms = dict(predefined or {})
    image = bytearray([fill] * size)

    def run(strict):
        pc = origin
        for line in lines:
            s = line.strip()
            m = re.match(r"^([A-Za-z_][A-Za-z0-9_]*):\s*(.*)$", s)
            if m:
                syms[m.group(1)] = pc
                s = m.group(2).strip()
                if not s:
                    continue
            m = re.match(r"^([A-Za-z_][A-Za-z0-9_]*)\s*=\s*(.+)$", s)
            if m:
                v = _expr(m.group(2), syms, pc, strict)
                if v is not None:
                    syms[m.group(1)] = v
                continue
            parts = s.split(None, 1)
            op = parts[0].upper()
            arg = parts[1].strip() if len(parts) > 1 else ""
            if op == ".ORG":
                pc = _expr(arg, syms, pc, True)
                continue
            if op == ".ALIGN":
                n = _expr(arg, syms, pc, True)
                pc = (pc + n - 1) // n * n
                continue
            if op == ".BYTE":
                for a in _split_args(arg):
                    v = _expr(a, syms, pc, strict)
                    if strict:
                        if not -128 <= v <= 255:
                            raise AsmError("byte out of range: %s" % line)
                        image[pc - origin] = v & 0xFF
                    pc += 1
                continue
            if op == ".WORD":
                for a in _split_args(arg):
                    v = _expr(a, syms, pc, strict)
                    if strict:
                        image[pc - origin] = v & 0xFF
                        image[pc - origin + 1] = (v >> 8) & 0xFF
                    pc += 2
                continue
            if op == ".FILL":
                n, v = [_expr(a, syms, pc, True) for a in _split_args(arg)]
                if strict:
                    for i in range(n):
                        image[pc - origin + i] = v & 0xFF
                pc += n
                continue
            if op in _IMPLIED and not arg:
                if strict:
                    image[pc - origin] = _IMPLIED[op]
                pc += 1
                continue
            if op in _BRANCH:
                if strict:
                    t = _expr(arg, syms, pc, True)
                    d = t - (pc + 2)
                    if not -128 <= d <= 127:
                        raise AsmError("branch out of range: %s (%d)" % (line, d))
                    image[pc - origin] = _BRANCH[op]
                    image[pc - origin + 1] = d & 0xFF
                pc += 2
                continue
            if op not in _OPS:
                raise AsmError("unknown mnemonic: %s" % line)
            modes = _OPS[op]
            a = arg
            au = a.upper().replace(" ", "")
            if au in ("", "A") and "acc" in modes:
                mode, e = "acc", None
            elif a.startswith("#"):
                mode, e = "imm", a[1:]
            elif re.match(r"^\(.*,X\)$", au):
                mode, e = "inx", a[a.index("(") + 1:a.upper().rindex(",X")]
            elif re.match(r"^\(.*\),Y$", au):
                mode, e = "iny", a[a.index("(") + 1:a.rindex(")")]
            elif re.match(r"^\(.*\)$", au) and "ind" in modes:
                mode, e = "ind", a[1:-1]
            elif au.endswith(",X"):
                mode, e = "?x", a[:a.upper().rindex(",X")]
            elif au.endswith(",Y"):
                mode, e = "?y", a[:a.upper().rindex(",Y")]
            else:
                mode, e = "?", a
            v = _expr(e, syms, pc, strict) if e is not None else None
            if mode in ("?", "?x", "?y"):
                suffix = {"?": "", "?x": "x", "?y": "y"}[mode]
                zpm, abm = "zp" + suffix, "ab" + (suffix or "s")
                if suffix == "":
                    abm = "abs"
                force_abs = e.strip().startswith("a:")
                if v is not None and 0 <= v < 0x100 and zpm in modes and not force_abs:
                    mode = zpm
                elif abm in modes:
                    mode = abm
                elif zpm in modes:
                    mode = zpm
                else:
                    raise AsmError("no such addressing mode: %s" % line)
            if mode not in modes:
                raise AsmError("bad addressing mode %s: %s" % (mode, line))
            n = _LEN[mode]
            if strict:
                image[pc - origin] = modes[mode]
                if n == 2:
                    if not -128 <= v <= 255:
                        raise AsmError("operand out of range: %s" % line)
                    image[pc - origin + 1] = v & 0xFF
                elif n == 3:
                    image[pc - origin + 1] = v & 0xFF
                    image[pc - origin + 2] = (v >> 8) & 0xFF
            pc += n
        return pc

    run(False)
    run(False)      # second sizing pass settles forward references used in `=` definitions
    run(True)
    return bytes(image), syms
