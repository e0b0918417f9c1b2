"""Randomised differential check of the scene loader's JSON reader (ovis_json_canonical) against the rapidjson the
reference vendors and parses its worlds with (oracle/json_harness.cc; json_serialiser.cc:213-214): generated documents
(nesting, every number spelling, escapes, whitespace) must give the same canonical text, and damaged ones (bytes dropped,
swapped, inserted) must be accepted or refused alike. Runs on the CPU (no device work).

    python tools/fuzz_json.py [--cases 20000] [--seed 1]"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from orion_b200 import lib  # noqa: E402


def both(so, ref, text: bytes):
    n = C.c_int64()
    buf = C.create_string_buffer(64 + 16 * len(text))
    rc = so.ovis_json_canonical(text, C.c_int64(len(text)), buf, C.c_int64(len(buf)), C.byref(n))
    ours = buf.raw[:n.value] if rc == lib.OVIS_OK else None
    rbuf = C.create_string_buffer(64 + 16 * len(text))
    m = ref.oref_json_canonical(text, C.c_int64(len(text)), rbuf, C.c_int64(len(rbuf)))
    theirs = rbuf.raw[:m] if m >= 0 else None
    return ours, theirs


def number(rng):
    k = rng.integers(0, 9)
    if k == 0:
        return str(int(rng.integers(-1000, 1000)))
    if k == 1:
        return str(int(rng.integers(-2**63, 2**63 - 1)))
    if k == 2:
        return repr(float(rng.normal() * 10.0 ** rng.integers(-40, 40)))
    if k == 3:
        return "%d.%0*d" % (rng.integers(-50, 50), int(rng.integers(1, 25)), rng.integers(0, 10**9))
    if k == 4:
        return "%de%s%d" % (rng.integers(-99, 99), rng.choice(["", "+", "-"]), rng.integers(0, 60))
    if k == 5:
        return "%d.%dE%s%d" % (rng.integers(0, 9), rng.integers(0, 99999), rng.choice(["", "+", "-"]), rng.integers(0, 400))
    if k == 6:
        return rng.choice(["0", "-0", "0.0", "-0.0", "0e0", "1e-400", "1e400", "-1e400", "16777217", "16777216.5", "0.1", "3.4028235e38",
                           "3.4028236e38", "1.401298464324817e-45", "7.006492321624085e-46", "340282356779733661637539395458142568448",
                           "18446744073709551615", "18446744073709551616", "-9223372036854775808", "-9223372036854775809",
                           "4294967295", "4294967296", "-2147483648", "-2147483649", "0.30000000000000004", "5e-324"])
    if k == 7:
        return "".join(rng.choice(list("0123456789"), int(rng.integers(1, 40))))             # leading zeros: both must refuse
    return repr(float(np.float32(rng.normal() * 100)))


def string(rng):
    parts = []
    for _ in range(int(rng.integers(0, 8))):
        k = rng.integers(0, 6)
        if k == 0:
            parts.append(rng.choice(['\\"', "\\\\", "\\/", "\\b", "\\f", "\\n", "\\r", "\\t"]))
        elif k == 1:
            parts.append("\\u%04x" % rng.integers(0, 0xD800))
        elif k == 2:
            parts.append("\\ud83d\\ude00")                                                   # a surrogate pair
        elif k == 3:
            parts.append(rng.choice(["é", "€", "日本", "\x7f"]))
        else:
            parts.append("".join(rng.choice(list("abcXYZ_ 09.-"), int(rng.integers(0, 6)))))
    return '"' + "".join(parts) + '"'


def value(rng, depth):
    ws = lambda: "".join(rng.choice([" ", "\n", "\t", "\r", ""], int(rng.integers(0, 3))))   # noqa: E731
    k = rng.integers(0, 8) if depth < 5 else rng.integers(0, 5)
    if k <= 1:
        return number(rng)
    if k == 2:
        return string(rng)
    if k == 3:
        return str(rng.choice(["true", "false", "null"]))
    if k == 4:
        return number(rng)
    if k <= 6:
        items = [ws() + value(rng, depth + 1) + ws() for _ in range(int(rng.integers(0, 5)))]
        return "[" + ",".join(items) + "]" if items else "[" + ws() + "]"
    members = [ws() + string(rng) + ws() + ":" + ws() + value(rng, depth + 1) + ws() for _ in range(int(rng.integers(0, 5)))]
    return "{" + ",".join(members) + "}" if members else "{" + ws() + "}"


def damage(rng, text: bytes) -> bytes:
    b = bytearray(text)
    for _ in range(int(rng.integers(1, 4))):
        if not b:
            break
        i = int(rng.integers(0, len(b)))
        k = rng.integers(0, 4)
        if k == 0:
            del b[i]
        elif k == 1:
            b.insert(i, int(rng.choice(list(b'{}[]",:0-+eE. \\tnu\x00\xff'))))
        elif k == 2:
            b[i] = int(rng.integers(0, 256))
        else:
            del b[i:]
    return bytes(b)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=20000)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    from oracle import pyoracle
    o = pyoracle.load("reference")
    ref = o.lib
    ref.oref_json_canonical.restype = C.c_int64
    so = lib.load()
    rng = np.random.default_rng(a.seed)
    accepted = refused = 0
    for case in range(a.cases):
        text = value(rng, 0).encode("utf-8", "surrogatepass")
        if rng.random() < 0.15:                                # rapidjson's documents: any value at the root
            text = b" " + text + b"\n"
        for t in (text, damage(rng, text)):
            ours, theirs = both(so, ref, t)
            if ours != theirs:
                print(json.dumps({"case": case, "text": t.decode("latin-1"), "ours": None if ours is None else ours.decode("latin-1"),
                                  "theirs": None if theirs is None else theirs.decode("latin-1")}))
                sys.exit(1)
            if ours is None:
                refused += 1
            else:
                accepted += 1
    print(json.dumps({"documents": 2 * a.cases, "accepted_by_both": accepted, "refused_by_both": refused, "differences": 0}))


if __name__ == "__main__":
    main()
