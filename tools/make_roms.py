"""Build the synthetic cartridges + their game descriptors (softtreemax_b200/roms/).

No Atari ROM can be obtained offline, and the ones the reference trains on are copyrighted
(reference Dockerfile:13-17 downloads them).  The stand-ins below are small original games written
for this repo; they keep what matters to the hot path: the ALE minimal-action-set size of the game
they stand for (4 / 6 / 9 actions -> 16/64, 216, 81 leaves), a controller of the same kind
(paddle / joystick), BCD score bytes, a lives byte, a real display kernel (VSYNC, VBLANK timer,
192 visible lines with per-line TIA writes, overscan) and collision-driven game logic.

    python tools/make_roms.py        # rewrites softtreemax_b200/roms/*.bin and games.json
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(__file__))
from asm6502 import assemble  # noqa: E402

TIA = """
VSYNC = $00
VBLANK = $01
WSYNC = $02
NUSIZ0 = $04
NUSIZ1 = $05
COLUP0 = $06
COLUP1 = $07
COLUPF = $08
COLUBK = $09
CTRLPF = $0A
REFP0 = $0B
REFP1 = $0C
PF0 = $0D
PF1 = $0E
PF2 = $0F
RESP0 = $10
RESP1 = $11
RESM0 = $12
RESM1 = $13
RESBL = $14
GRP0 = $1B
GRP1 = $1C
ENAM0 = $1D
ENAM1 = $1E
ENABL = $1F
HMP0 = $20
HMP1 = $21
HMM0 = $22
HMM1 = $23
HMBL = $24
VDELP0 = $25
VDELP1 = $26
VDELBL = $27
RESMP0 = $28
RESMP1 = $29
HMOVE = $2A
HMCLR = $2B
CXCLR = $2C
CXM0P = $00
CXM1P = $01
CXP0FB = $02
CXP1FB = $03
CXM0FB = $04
CXM1FB = $05
CXBLPF = $06
CXPPMM = $07
INPT0 = $08
INPT4 = $0C
SWCHA = $0280
SWCHB = $0282
INTIM = $0284
TIM64T = $0296
"""

_uid = [0]


def line_body(extra=""):
    """One visible scanline: ball enable by vertical range, paddle-pot sampling, optional extra."""
    _uid[0] += 1
    u = _uid[0]
    return f"""
    STA WSYNC
    {extra}
    TYA
    SEC
    SBC ballY
    CMP #BALL_H
    LDA #0
    BCS nb{u}
    LDA #2
nb{u}: STA ENABL
    BIT INPT0
    BMI ch{u}
    INC padCnt
ch{u}:
"""


def region(name, end, extra=""):
    return f"""
{name}:
{line_body(extra)}
    INY
    CPY #{end}
    BNE {name}
"""


def breakout_source(variant):
    """variant: 'breakout' (paddle, lives), 'pong' (joystick, opponent score), 'pacman' (F8)."""
    paddle = variant == "breakout"
    banked = variant == "pacman"
    src = TIA + """
frameCnt = $80
ballX = $81
ballY = $82
ballDX = $83
ballDY = $84
padX = $85
padCnt = $86
ballLive = $87
temp = $88
temp2 = $89
bricks = $90
oppScore = $B8
lives = $B9
gameOver = $BA
scoreHi = $CC
scoreLo = $CD
padSize = $BB

BALL_H = 4
BRICK_TOP = 32
PADDLE_Y = 176

    .org $F000
Reset:
    SEI
    CLD
    LDX #$FF
    TXS
    LDA #0
    TAX
clr: STA 0,X
    INX
    BNE clr
    LDA #5
    STA lives
    LDA #76
    STA padX
    JSR FillBricks
    LDA #$21            ; reflect playfield, ball 4 clocks wide
    STA CTRLPF
    LDA #5              ; double-width paddle
    STA NUSIZ0
    LDA #$0E
    STA COLUPF
    LDA #$46
    STA COLUP0
    LDA #$00
    STA COLUBK

MainLoop:
    LDA #2
    STA VSYNC           ; <- frame boundary for the emulator
    STA WSYNC
    STA WSYNC
    STA WSYNC
    LDA #0
    STA VSYNC
    LDA #43
    STA TIM64T
    LDA #$82
    STA VBLANK
    INC frameCnt
    JSR GameLogic
""" + ("    JSR CallBank1\n" if banked else "") + """
    ; position paddle (P0) and ball (BL)
    LDA padX
    LDX #0
    JSR PosObject
    LDA ballX
    LDX #4
    JSR PosObject
    STA WSYNC
    STA HMOVE
waitVB: LDA INTIM
    BNE waitVB
    LDA #0
    STA padCnt
    TAY                 ; Y = visible line counter
    STA WSYNC
    STA HMCLR
    STA VBLANK          ; display on, paddle caps start charging

    ; ---- visible kernel, 192 lines in regions ----
    LDA #$F0
    STA PF0
    LDA #$FF
    STA PF1
    STA PF2
""" + region("topWall", 8) + """
    LDA #$10
    STA PF0
    LDA #0
    STA PF1
    STA PF2
""" + region("upper", "BRICK_TOP") + """
    LDX #0
brickRow:
    LDA bricks,X
    STA temp
    LDA bricks+1,X
    STA temp2
    STX padX+64         ; scratch save of X (RAM $C5)
    TYA
    CLC
    ADC #8
    STA padX+65         ; end line of this row (RAM $C6)
brickLine:
""" + line_body("LDA temp\n    STA PF1\n    LDA temp2\n    STA PF2") + """
    INY
    CPY padX+65
    BNE brickLine
    LDX padX+64
    INX
    INX
    CPX #12
    BNE brickRow
    LDA #0
    STA PF1
    STA PF2
""" + region("middle", "PADDLE_Y") + """
    LDA #$FF
    STA GRP0
""" + region("padRows", "PADDLE_Y+8") + """
    LDA #0
    STA GRP0
""" + region("bottom", 192) + """
    ; ---- overscan ----
    STA WSYNC
    LDA #$82
    STA VBLANK
    LDA #0
    STA ENABL
    STA PF0
    LDA #34
    STA TIM64T
waitOS: LDA INTIM
    BNE waitOS
    JMP MainLoop

; ------------------------------------------------------------------
; A = x (0..159), X = object (0 P0, 1 P1, 2 M0, 3 M1, 4 BL)
PosObject:
    STA WSYNC
    SEC
divl: SBC #15
    BCS divl
    EOR #7
    ASL
    ASL
    ASL
    ASL
    STA HMP0,X
    STA RESP0,X
    RTS

FillBricks:
    LDX #11
    LDA #$FF
fb1: STA bricks,X
    DEX
    BPL fb1
    RTS

; ------------------------------------------------------------------
GameLogic:
    LDA gameOver
    BEQ alive
    RTS
alive:
"""
    if paddle:
        src += """
    ; paddle x from the pot reading of the previous frame
    LDA #168
    SEC
    SBC padCnt
    BCS px1
    LDA #4
px1: CMP #4
    BCS px2
    LDA #4
px2: CMP #141
    BCC px3
    LDA #140
px3: STA padX
    LDA SWCHA           ; paddle 0 button = bit 7, low when pressed
    STA temp
"""
    else:
        src += """
    ; joystick: left/right move the paddle 3 px per frame; fire = INPT4
    LDA SWCHA
    ASL                 ; C = right bit (0 = pushed)
    BCS noRight
    LDA padX
    CLC
    ADC #3
    CMP #141
    BCC pr1
    LDA #140
pr1: STA padX
noRight:
    LDA SWCHA
    AND #$40
    BNE noLeft
    LDA padX
    SEC
    SBC #3
    CMP #4
    BCS pl1
    LDA #4
pl1: STA padX
noLeft:
""" + ("""
    LDA SWCHA           ; no FIRE in this action set: any stick direction launches the ball
    AND #$F0
    CMP #$F0
    BEQ noDir
    LDA #0
    BEQ stDir
noDir: LDA #$80
stDir: STA temp
""" if banked else """
    LDA INPT4
    STA temp
""")
    src += """
    LDA ballLive
    BNE inPlay
    ; ball rides on the paddle until FIRE
    LDA padX
    CLC
    ADC #7
    STA ballX
    LDA #PADDLE_Y-6
    STA ballY
    BIT temp
    BMI glDone0
    LDA #1
    STA ballLive
    STA ballDX
    LDA #$FE
    STA ballDY
glDone0:
    STA CXCLR
    RTS

inPlay:
    ; --- paddle hit ---
    BIT CXP0FB
    BVC noPad
    LDA ballDY
    BMI noPad           ; already going up
    LDA #$FE
    STA ballDY
    LDA ballX
    SEC
    SBC padX
    CMP #8
    BCS padRight
    LDA #$FF
    STA ballDX
    BNE noPad
padRight:
    LDA #1
    STA ballDX
noPad:
    ; --- brick hit ---
    BIT CXBLPF
    BPL noBrick
    LDA ballY
    LDX ballDY
    BMI probeTop
    CLC
    ADC #BALL_H-1
probeTop:
    SEC
    SBC #BRICK_TOP
    CMP #48
    BCS noBrick
    LSR
    LSR
    LSR
    STA temp2           ; row 0..5
    LDA ballX
    CMP #80
    BCC leftHalf
    LDA #159
    SEC
    SBC ballX
leftHalf:
    LSR
    LSR                 ; playfield bit index 0..19
    CMP #4
    BCC noBrick
    SEC
    SBC #4
    TAY                 ; 0..15
    LDA temp2
    ASL
    TAX                 ; 2*row
    CPY #8
    BCC inPF1
    INX
inPF1:
    LDA bricks,X
    AND BitTab,Y
    BEQ noBrick
    LDA BitTab,Y
    EOR #$FF
    AND bricks,X
    STA bricks,X
    ; bounce + score
    LDA #0
    SEC
    SBC ballDY
    STA ballDY
    LDX temp2
    SED
    CLC
    LDA scoreLo
    ADC RowScore,X
    STA scoreLo
    LDA scoreHi
    ADC #0
    STA scoreHi
    CLD
    ; refill when the wall is gone
    LDX #11
    LDA #0
chk: ORA bricks,X
    DEX
    BPL chk
    CMP #0
    BNE noBrick
    JSR FillBricks
noBrick:
    STA CXCLR
    ; --- move ---
    LDA ballX
    CLC
    ADC ballDX
    STA ballX
    CMP #5
    BCS xlo
    LDA #1
    STA ballDX
xlo: LDA ballX
    CMP #151
    BCC xhi
    LDA #$FF
    STA ballDX
xhi: LDA ballY
    CLC
    ADC ballDY
    STA ballY
    CMP #9
    BCS ylo
    LDA #2
    STA ballDY
ylo: LDA ballY
    CMP #188
    BCC glDone
    ; ball lost
    LDA #0
    STA ballLive
"""
    if variant == "pong":
        src += """
    SED
    CLC
    LDA oppScore
    ADC #1
    STA oppScore
    CLD
    CMP #$21
    BNE glDone
    LDA #1
    STA gameOver
"""
    else:
        src += """
    DEC lives
    BNE glDone
    LDA #1
    STA gameOver
"""
    src += """
glDone:
    RTS

BitTab:
    .byte $80,$40,$20,$10,$08,$04,$02,$01
    .byte $01,$02,$04,$08,$10,$20,$40,$80
RowScore:
    .byte $07,$07,$04,$04,$01,$01
"""
    if banked:
        src += """
; ---- F8 bank switching: these addresses exist in both banks ----
    .org $FFD0
CallBank1:
    LDA $1FF9           ; -> bank 1: the next fetch ($FFD3) comes from bank 1 (JSR Bank1Logic ; LDA $1FF8)
    NOP
    NOP
    NOP
    NOP
    NOP
    NOP
    LDA padSize         ; $FFD9: bank 1 switched back to bank 0 just before this address
    STA NUSIZ0
    RTS
"""
    src += """
    .org $FFFA
    .word Reset, Reset, Reset
"""
    return src


def torture_source():
    """TIA/RIOT torture cart: every visible line writes LFSR-chosen values to LFSR-chosen TIA
    registers at LFSR-chosen horizontal times; collisions feed the BCD score; joystick and fire
    are folded into the LFSR so that sibling actions diverge.  Exists for parity coverage."""
    regs = ["NUSIZ0", "NUSIZ1", "COLUP0", "COLUP1", "COLUPF", "COLUBK", "CTRLPF", "REFP0", "REFP1",
            "PF0", "PF1", "PF2", "RESP0", "RESP1", "RESM0", "RESM1", "RESBL", "GRP0", "GRP1", "ENAM0",
            "ENAM1", "ENABL", "HMP0", "HMP1", "HMM0", "HMM1", "HMBL", "VDELP0", "VDELP1", "VDELBL",
            "RESMP0", "RESMP1", "HMOVE", "HMCLR",
            "GRP0", "GRP1", "ENAM0", "ENAM1", "ENABL", "GRP0", "GRP1", "PF1", "PF2", "CTRLPF",
            "NUSIZ0", "NUSIZ1", "RESP0", "RESP1", "RESBL", "COLUP0", "COLUP1", "COLUPF", "HMOVE",
            "GRP0", "GRP1", "ENAM0", "ENAM1", "ENABL", "REFP0", "REFP1", "RESM0", "RESM1", "PF0", "COLUBK"]
    assert len(regs) == 64
    return TIA + """
lfsr = $80
cnt = $81
acc = $82
ptr = $84
scoreHi = $CC
scoreLo = $CD
    .org $F000
Reset:
    SEI
    CLD
    LDX #$FF
    TXS
    LDA #0
    TAX
clr: STA 0,X
    INX
    BNE clr
    LDA #$A5
    STA lfsr
    LDA #<RegTab
    STA ptr
    LDA #>RegTab
    STA ptr+1
Main:
    LDA #2
    STA VSYNC
    STA WSYNC
    STA WSYNC
    STA WSYNC
    LDA #0
    STA VSYNC
    LDA #43
    STA TIM64T
    LDA #$42
    STA VBLANK
    ; fold input into the LFSR
    LDA SWCHA
    EOR lfsr
    ADC INPT4
    STA lfsr
    ; collisions of the last frame -> score
    LDA #0
    LDX #7
cxl: ORA CXM0P,X
    DEX
    BPL cxl
    STA CXCLR
    STA acc
    ASL
    BCC nocx7
    SED
    CLC
    LDA scoreLo
    ADC #1
    STA scoreLo
    LDA scoreHi
    ADC #0
    STA scoreHi
    CLD
nocx7:
    BIT acc
    BVC nocx6
    PHP
    SED
    SEC
    LDA scoreLo
    SBC #0
    PLP
nocx6:
waitVB: LDA INTIM
    BNE waitVB
    STA WSYNC
    STA VBLANK
    LDY #192
line:
    STA WSYNC
    JSR Rand
    AND #7
    TAX
dl1: DEX
    BPL dl1
    JSR Rand
    AND #$3F
    STY cnt
    TAY
    LDA (ptr),Y
    TAX
    LDY cnt
    JSR Rand
    STA 0,X
    JSR Rand
    AND #3
    TAX
dl2: DEX
    BPL dl2
    JSR Rand
    AND #$3F
    TAX
    LDA RegTab,X
    TAX
    LDA lfsr
    EOR cnt
    STA 0,X
    DEY
    BNE line
    LDA #2
    STA VBLANK
    LDA #30
    STA TIM64T
waitOS: LDA INTIM
    BNE waitOS
    JMP Main

Rand:
    LDA lfsr
    LSR
    BCC nr
    EOR #$B8
nr: STA lfsr
    RTS

RegTab:
    .byte """ + ", ".join(regs) + """
    .org $FFFA
    .word Reset, Reset, Reset
"""


def pacman_bank1_source():
    """Bank 1 of the F8 cartridge: entered through the trampoline at $FE00; joystick UP / DOWN pick the
    paddle width (quad / single), anything else restores double width."""
    return TIA + """
padSize = $BB
    .org $F000
Bank1Logic:
    LDA SWCHA
    AND #$10            ; up (0 = pushed)
    BNE notUp
    LDA #7
    STA padSize
    RTS
notUp:
    LDA SWCHA
    AND #$20            ; down
    BNE notDown
    LDA #0
    STA padSize
    RTS
notDown:
    LDA #5
    STA padSize
    RTS

    .org $FFD0
    .byte $EA, $EA, $EA ; (bank 0 executes LDA $1FF9 here)
    JSR Bank1Logic      ; at $FFD3 in bank 1
    LDA $1FF8           ; -> bank 0, which continues at $FFD9
    .org $FFFA
    .word $F000, $F000, $F000
"""


ALE = dict(NOOP=0, FIRE=1, UP=2, RIGHT=3, LEFT=4, DOWN=5, UPRIGHT=6, UPLEFT=7, DOWNRIGHT=8,
           DOWNLEFT=9, UPFIRE=10, RIGHTFIRE=11, LEFTFIRE=12, DOWNFIRE=13)


def build():
    out = os.path.join(os.path.dirname(__file__), "..", "softtreemax_b200", "roms")
    os.makedirs(out, exist_ok=True)
    games = {}

    rom, _ = assemble(breakout_source("breakout"), origin=0xF000, size=4096)
    open(os.path.join(out, "synth_breakout.bin"), "wb").write(rom)
    games["synth_breakout"] = dict(
        file="synth_breakout.bin", stands_for="Breakout", bank_scheme=0, controller=1,
        actions=[ALE["NOOP"], ALE["FIRE"], ALE["RIGHT"], ALE["LEFT"]],
        score_addr=[255, 76, 77], score2_addr=[255, 255, 255],
        lives_addr=57, lives_shift=0, lives_mask=255, term_addr=57, term_mask=255, term_value=0)

    rom, _ = assemble(breakout_source("pong"), origin=0xF000, size=4096)
    open(os.path.join(out, "synth_pong.bin"), "wb").write(rom)
    games["synth_pong"] = dict(
        file="synth_pong.bin", stands_for="Pong", bank_scheme=0, controller=0,
        actions=[ALE["NOOP"], ALE["FIRE"], ALE["RIGHT"], ALE["LEFT"], ALE["RIGHTFIRE"], ALE["LEFTFIRE"]],
        score_addr=[255, 76, 77], score2_addr=[255, 255, 56],
        lives_addr=255, lives_shift=0, lives_mask=255, term_addr=58, term_mask=255, term_value=1)

    bank0, _ = assemble(breakout_source("pacman"), origin=0xF000, size=4096)
    bank1, _ = assemble(pacman_bank1_source(), origin=0xF000, size=4096)
    bank1 = bytearray(bank1)
    bank1[0xFFC:0x1000] = bank0[0xFFC:0x1000]           # both banks carry the reset vector of bank 0
    open(os.path.join(out, "synth_pacman.bin"), "wb").write(bank0 + bytes(bank1))
    games["synth_pacman"] = dict(
        file="synth_pacman.bin", stands_for="MsPacman", bank_scheme=1, controller=0,
        actions=[ALE["NOOP"], ALE["UP"], ALE["RIGHT"], ALE["LEFT"], ALE["DOWN"], ALE["UPRIGHT"], ALE["UPLEFT"],
                 ALE["DOWNRIGHT"], ALE["DOWNLEFT"]],
        score_addr=[255, 76, 77], score2_addr=[255, 255, 255],
        lives_addr=57, lives_shift=0, lives_mask=255, term_addr=57, term_mask=255, term_value=0)

    rom, _ = assemble(torture_source(), origin=0xF000, size=4096)
    open(os.path.join(out, "synth_torture.bin"), "wb").write(rom)
    games["synth_torture"] = dict(
        file="synth_torture.bin", stands_for="(parity coverage)", bank_scheme=0, controller=0,
        actions=list(range(18)),
        score_addr=[255, 76, 77], score2_addr=[255, 255, 255],
        lives_addr=255, lives_shift=0, lives_mask=255, term_addr=255, term_mask=255, term_value=0)

    # Action-count classes of the reference's sweep (yamls/example_sweep.yml:19-26: Asteroids 14, Breakout 4, Gopher 8, Krull 18,
    # KungFuMaster 14, NameThisGame 6, Phoenix 8, VideoPinball 9 minimal actions): the joystick cartridge under wider action
    # sets.  What a tree costs depends on the branching factor A (A^d leaves per root), not on which game draws the frames.
    fire = lambda names: [ALE[n] for n in names]
    base = dict(games["synth_pong"])
    games["synth_joy8"] = dict(base, stands_for="Gopher / Phoenix (8 actions)",
                               actions=fire(["NOOP", "FIRE", "UP", "RIGHT", "LEFT", "UPFIRE", "RIGHTFIRE", "LEFTFIRE"]))
    games["synth_joy14"] = dict(base, stands_for="Asteroids / KungFuMaster (14 actions)",
                                actions=fire(["NOOP", "FIRE", "UP", "RIGHT", "LEFT", "DOWN", "UPRIGHT", "UPLEFT", "UPFIRE", "RIGHTFIRE",
                                              "LEFTFIRE", "DOWNFIRE"]) + [14, 15])
    games["synth_joy18"] = dict(base, stands_for="Krull (18 actions)", actions=list(range(18)))
    json.dump(games, open(os.path.join(out, "games.json"), "w"), indent=1, sort_keys=True)
    return games


if __name__ == "__main__":
    g = build()
    print("built", ", ".join(sorted(g)))
