"""Extracts, from Rust-compiled extension modules that happen to be installed in this image, the machine code of
fastrand's WyRand step — a compiled copy of the THIRD-PARTY dependency the reference draws every random number from
(fastrand 2.3.0, the reference's Cargo.lock:338-341; call sites math/vector.rs:31-33, 37-38).  The reference cannot be
compiled here (no cargo/rustc) and ships no known answers for its RNG, so until round 2 the oracle's generator rested
on recollection; this pins it to an artifact: hf_xet 1.5.0 ships a CycloneDX SBOM that lists `pkg:cargo/fastrand@2.3.0`,
and its .so holds `fastrand::Rng::gen_u64` inlined (symbols `fastrand::global_rng::RNG` are in the same file).

Writes tests/golden/fastrand_wyrand_step.json.  tests/test_fastrand_pin.py executes the extracted instructions with a
small x86-64 interpreter and compares the stream with the oracle's.  Run from the repo root:
  python tests/golden/extract_fastrand_evidence.py
"""
import glob
import json
import mmap
import os
import struct
import sysconfig

C_ADD = struct.pack("<Q", 0x2D358DCCAA6C78A5)  # state += ...
C_XOR = struct.pack("<Q", 0x8BB84B93962EACC9)  # t = state ^ ...
HERE = os.path.dirname(os.path.abspath(__file__))


def sbom_fastrand_version(dist_info):
    for f in glob.glob(os.path.join(dist_info, "sboms", "*.json")):
        try:
            d = json.load(open(f))
        except Exception:
            continue
        for c in d.get("components", []):
            if c.get("name") == "fastrand":
                return c.get("version"), c.get("purl"), os.path.basename(f)
    return None, None, None


def find(site):
    out = []
    for so in glob.glob(os.path.join(site, "*", "*.so")):
        if os.path.getsize(so) < 100_000:
            continue
        with open(so, "rb") as fh:
            m = mmap.mmap(fh.fileno(), 0, access=mmap.ACCESS_READ)
            a = m.find(C_XOR)
            # `movabs reg, C_XOR` (REX.W+B8+r, 10 bytes) directly followed by `movabs reg, C_ADD`
            if a >= 2 and m[a + 8 + 2:a + 8 + 10] == C_ADD and m.find(b"fastrand") >= 0:
                pkg = os.path.basename(os.path.dirname(so))
                di = glob.glob(os.path.join(site, pkg.replace("-", "_") + "-*.dist-info"))
                ver = os.path.basename(di[0]).split("-")[1].replace(".dist", "") if di else None
                fv, purl, sbom = sbom_fastrand_version(di[0]) if di else (None, None, None)
                out.append({"package": pkg, "package_version": ver, "file": os.path.relpath(so, site), "offset": a - 2,
                            "sbom": sbom, "sbom_fastrand_version": fv, "sbom_purl": purl,
                            "code_hex": bytes(m[a - 2:a - 2 + 72]).hex()})
            m.close()
    return out


if __name__ == "__main__":
    site = sysconfig.get_paths()["purelib"]
    found = find(site)
    if not found:
        raise SystemExit("no installed Rust extension module with an inlined fastrand WyRand step was found")
    found.sort(key=lambda e: (e["sbom_fastrand_version"] is None, e["package"]))
    json.dump({"what": "machine code of fastrand's WyRand step as compiled into third-party wheels of this image",
               "reference_pin": "fastrand 2.3.0 (Cargo.lock:338-341 of the reference)", "entries": found},
              open(os.path.join(HERE, "fastrand_wyrand_step.json"), "w"), indent=1)
    for e in found:
        print(e["package"], e["package_version"], "fastrand", e["sbom_fastrand_version"], e["code_hex"][:60], "...")
