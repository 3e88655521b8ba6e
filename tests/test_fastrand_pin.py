"""The oracle's random source against a COMPILED copy of the reference's own dependency.

The reference draws every random number from fastrand 2.3.0 (Cargo.lock:338-341; call sites math/vector.rs:31-33,
37-38).  It cannot be built here and holds no known answers for it, so the oracle's WyRand (oracle/rtx_oracle.cpp,
`wy_u64`; rust-raytracing_b200/csrc/rtx_rng.cuh on the device) used to rest on recollection.  Two Rust wheels installed
in this image carry fastrand with `Rng::gen_u64` inlined; one of them (hf_xet 1.5.0) ships an SBOM that says
`pkg:cargo/fastrand@2.3.0` — the reference's exact pin.  tests/golden/extract_fastrand_evidence.py cut the machine code
of that step out of the .so files (tests/golden/fastrand_wyrand_step.json); here a small x86-64 interpreter EXECUTES those
instructions and the stream they produce must be the oracle's, bit for bit.  (What stays unpinned is the f64 mantissa
fill of `Rng::f64` — none of the wheels calls it — and everything that is not the RNG.)"""
import json
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(__file__), "golden", "fastrand_wyrand_step.json")
M64 = (1 << 64) - 1
REGS = ["rax", "rcx", "rdx", "rbx", "rsp", "rbp", "rsi", "rdi", "r8", "r9", "r10", "r11", "r12", "r13", "r14", "r15"]


class X86:
    """mov r64, imm64 | mov r64, r/m64 | mov r/m64, r64 | add/xor r/m64, r64 | mul r/m64 | lea r64, [base + index*scale]
    — 64-bit operand size only, register or [reg] / [base+index*scale] operands without displacement."""

    def __init__(self, code):
        self.code, self.ip, self.r, self.mem, self.trace = code, 0, {n: 0 for n in REGS}, {}, []

    def _modrm(self, rex):
        b = self.code[self.ip]
        self.ip += 1
        mod, reg, rm = b >> 6, ((b >> 3) & 7) | ((rex >> 2) & 1) << 3, b & 7
        if mod == 3:
            return reg, ("r", REGS[rm | (rex & 1) << 3])
        assert mod == 0, "displacements are not modelled"
        if rm == 4:  # SIB
            s = self.code[self.ip]
            self.ip += 1
            scale, idx, base = 1 << (s >> 6), ((s >> 3) & 7) | ((rex >> 1) & 1) << 3, (s & 7) | (rex & 1) << 3
            assert (s & 7) != 5 and idx != 4
            return reg, ("m", (self.r[REGS[base]] + self.r[REGS[idx]] * scale) & M64)
        assert rm != 5
        return reg, ("m", self.r[REGS[rm | (rex & 1) << 3]])

    def _get(self, op):
        return self.r[op[1]] if op[0] == "r" else self.mem[op[1]]

    def _set(self, op, v):
        if op[0] == "r":
            self.r[op[1]] = v & M64
        else:
            self.mem[op[1]] = v & M64

    def step(self):
        """Executes one instruction; False when the next one is outside the modelled subset."""
        ip0 = self.ip
        if self.ip + 3 > len(self.code):
            return False
        rex = self.code[self.ip]
        if not (0x48 <= rex <= 0x4F):  # REX.W required
            return False
        op = self.code[self.ip + 1]
        self.ip += 2
        if 0xB8 <= op <= 0xBF:
            dst = REGS[(op & 7) | (rex & 1) << 3]
            self.r[dst] = int.from_bytes(self.code[self.ip:self.ip + 8], "little")
            self.ip += 8
            self.trace.append(("mov_imm", dst))
        elif op in (0x89, 0x8B, 0x01, 0x31, 0x8D):
            reg, rm = self._modrm(rex)
            rn = REGS[reg]
            if op == 0x8B:
                self.r[rn] = self._get(rm)
            elif op == 0x89:
                self._set(rm, self.r[rn])
            elif op == 0x01:
                self._set(rm, self._get(rm) + self.r[rn])
            elif op == 0x31:
                self._set(rm, self._get(rm) ^ self.r[rn])
            else:
                assert rm[0] == "m"
                self.r[rn] = rm[1]
            self.trace.append(({0x89: "mov", 0x8B: "mov", 0x01: "add", 0x31: "xor", 0x8D: "lea"}[op], rn, rm))
        elif op == 0xF7 and ((self.code[self.ip] >> 3) & 7) == 4:
            _, rm = self._modrm(rex)
            p = self.r["rax"] * self._get(rm)
            self.r["rax"], self.r["rdx"] = p & M64, p >> 64
            self.trace.append(("mul", rm))
        else:
            self.ip = ip0
            return False
        return True


def _run_step(code, state):
    """One WyRand step of the compiled code: (output, new state).  The snippet is run up to and including the xor that
    follows its first mul (hi ^ lo); where the state lives is read off the code: [rbx] (hf_xet) or rdi -> rsi (pyqwest)."""
    cpu = X86(code)
    cpu.r["rbx"], cpu.mem[0x1000], cpu.r["rdi"] = 0x1000, state, state
    seen_mul = False
    while cpu.step():
        kind = cpu.trace[-1][0]
        if kind == "mul":
            seen_mul = True
        elif seen_mul and kind == "xor":
            out = cpu._get(cpu.trace[-1][2])
            new_state = cpu.mem[0x1000] if cpu.mem[0x1000] != state else cpu.r["rsi"]
            return out, new_state, cpu.trace
    raise AssertionError("the snippet did not reach mul + xor")


def test_compiled_fastrand_wyrand_step_is_the_oracles(oracle):
    from oracle import pyref
    entries = json.load(open(G))["entries"]
    assert any(e["sbom_fastrand_version"] == "2.3.0" for e in entries), "no SBOM-backed fastrand 2.3.0 entry"
    for e in entries:
        code = bytes.fromhex(e["code_hex"])
        for seed in (0, 1, 0x0EF6F79ED30BA75A, M64, 0x123456789ABCDEF0):
            state, got = seed, []
            for _ in range(8):
                out, state, trace = _run_step(code, state)
                got.append(out)
            kinds = [t[0] for t in trace]
            assert kinds.count("mul") == 1 and ("add" in kinds or "lea" in kinds), kinds
            want = oracle.wyrand_u64(seed, 8)[0]
            assert [int(v) for v in want] == got, (e["package"], hex(seed))
            w = pyref.WyRand(seed)
            assert [w.u64() for _ in range(8)] == got
            assert state == (seed + 8 * 0x2D358DCCAA6C78A5) & M64  # the state is a pure counter


def test_the_fixture_is_what_the_installed_wheels_hold():
    """Where the wheels are installed (this image, here and on the GPU box), the committed bytes are in their .so files
    and hf_xet's SBOM names fastrand 2.3.0; elsewhere this is skipped and the committed fixture stands alone."""
    import importlib.util
    import sysconfig
    site = sysconfig.get_paths()["purelib"]
    spec = importlib.util.spec_from_file_location("extract", os.path.join(os.path.dirname(G), "extract_fastrand_evidence.py"))
    ex = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ex)
    live = {e["package"]: e for e in ex.find(site)}
    if not live:
        pytest.skip("no Rust wheel with fastrand in this environment")
    committed = {e["package"]: e for e in json.load(open(G))["entries"]}
    for pkg, e in live.items():
        if pkg in committed and committed[pkg]["package_version"] == e["package_version"]:
            assert committed[pkg]["code_hex"] == e["code_hex"] and committed[pkg]["sbom_fastrand_version"] == e["sbom_fastrand_version"]
