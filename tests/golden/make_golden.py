"""Regenerates tests/golden/*.  Run in the build container (needs /root/reference
for the Criteo excerpt).  Committed so the provenance of every fixture is clear.

  parser_vectors.json   literal vectors of the reference's own parser test
                        (test/fm_parser_op_test.py:7-15, :48-81) + Hash64 known answers
                        (recalled from TF's hash_test.cc; the default-seed
                        values are self-generated regression values, NOT reference-pinned).
  scorer_case.json      the reference scorer test's fixed inputs (test/fm_scorer_op_test.py:10-23,
                        params drawn with a fixed seed) and the expected pred / reg / grad from
                        the test's own dense closed form + autograd (:49-63), in float64.
  strtof_vectors.json   decimal / hex / special strings with glibc strtof's bits and consumed
                        length (glibc IS the reference's arithmetic: fm_parser_op.cc:69,98).
  criteo_excerpt.txt / criteo_excerpt_weights.txt
                        first 2000 lines of the reference's data/train_0 and data/weight_0.
"""
import json
import os
import random
import struct
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import dense_ref, oracle as O  # noqa: E402


def parser_vectors():
    kat_seed = 0xbc9f1d34
    return {
        "examples": ["1 1:1 2:2 3:3 4:4", "-1 5:1 6:1 7:1 ", "1.0 8:0.1 9:0.2 10:0.3"],
        "vocab_size": 10000,
        "sizes": [4, 3, 3],
        "labels": [1, -1, 1.0],
        "feature_ids": [1, 2, 3, 4, 5, 6, 7, 8, 9, 10],
        "feature_vals": [1, 2, 3, 4, 1, 1, 1, 0.1, 0.2, 0.3],
        "errors": [
            ["one 1:1 2:2 3:3 4:4", "Label could not be read in example: "],
            ["1 one:1 2:2 3:3 4:4", "Invalid format in example: "],
            ["1 1:one 2:2 3:3 4:4", "Invalid feature value. "],
        ],
        "hash64_kat_seed": kat_seed,
        "hash64_kat": [
            ["", "4c61ea3eeda4cb87"], ["62", "091309f7ef916c8a"], ["c397", "a815bcdf1d1af01c"],
            ["e299a5", "02167564e4d06430"], ["e180b932", "8f7ed82ffc21071f"],
        ],
        "hash64_default_seed": 0xDECAFCAFFE,
        "hashed_ids_1_to_10_mod_10000_selfgenerated": [2333, 1362, 4535, 1205, 4140, 7902, 7642, 7211, 8813, 4343],
        "hash64_long_selfgenerated": [
            [s, "%016x" % O.hash64(s)] for s in
            ["12_0000abcd", "feature_name_longer_than_sixteen_bytes", "abcdefgh", "abcdefghi", "x" * 31]
        ],
    }


def scorer_case():
    rnd = random.Random(20161005)
    factor_num, vocabulary_size = 8, 10
    feature_ids = [[1, 2, 3, 2, 4, 9], [3, 4, 7, 0, 0, 0]]
    feature_vals = [[9.0, 2.0, 3.6, 4.0, -2.2, -10.7], [4.0, -3.4, 2.0, 0, 0, 0]]
    labels = [1.0, -1.0]
    params = [[0.0] * (factor_num + 1)]
    for _ in range(vocabulary_size):
        params.append([rnd.uniform(0.01, 0.02) for _ in range(factor_num + 1)])
    params = np.asarray(params, np.float32)
    flat_ids, flat_vals, poses = [], [], [0]
    for r_ids, r_vals in zip(feature_ids, feature_vals):
        for fid, fval in zip(r_ids, r_vals):
            if fid == 0:
                continue
            flat_ids.append(fid)
            flat_vals.append(fval)
        poses.append(len(flat_ids))
    pred, reg, grad = dense_ref.dense_cost_and_grad(
        flat_ids, torch.from_numpy(params), np.asarray(flat_vals, np.float32), poses, 1.2, 0.4, labels)
    return {
        "factor_num": factor_num, "factor_lambda": 1.2, "bias_lambda": 0.4, "labels": labels,
        "feature_ids": flat_ids, "feature_vals": flat_vals, "feature_poses": poses,
        "params": params.tolist(), "pred": pred.tolist(), "reg": float(reg), "grad": grad.tolist(),
        "note": "cost = sum((label-pred)^2) + reg; so pred_grad = -2(label-pred), reg_grad = 1",
    }


def strtof_vectors():
    rnd = random.Random(7)
    cases = ["0", "-0", "+0", "1", "-1", "1.0", "0.1", "0.2", "0.3", ".5", "5.", "-.5e1", "1e", "1e+", "1e-",
             "1e5", "1E5", "1e+05", "1e-05", "3.4028235e38", "3.4028236e38", "3.40282357e38", "3.5e38",
             "1e39", "-1e39", "1.1754944e-38", "1.17549435e-38", "1e-45", "1.4e-45", "7e-46", "7.1e-46",
             "0.7006492321624085e-45", "0.70064923216240854e-45", "1e-46", "1e-400", "1e400", "0e400",
             "16777216", "16777217", "16777218", "16777219", "33554434", "33554435", "9007199254740993",
             "0.000000000000000000000000000000000000000000001", "123456789012345678901234567890",
             "1.00000005960464477539062", "1.000000059604644775390625", "1.0000000596046447753906250000001",
             "1.00000017881393432617187", "1.000000178813934326171875", "1.0000001788139343261718750001",
             "8388608.5", "8388609.5", "8388608.50000000000000000000000000000001", "4.2e1x", "1.5.3", "1..2",
             "inf", "-inf", "+INF", "Infinity", "-INFINITY", "infinit", "in", "nan", "-nan", "NaN", "nan(", "nan()",
             "nan(123)", "nan(0x1f)", "nan(abc)", "-nan(7)", "nan(12", "nan(0)", "nan(4194304)", "nan(08)",
             "0x", "0x1", "0x1p3", "0x1.8p1", "0X.8P+1", "0x1.fffffep127", "0x1.ffffffp127", "0x1p-149", "0x1p-150",
             "0x1.000001p-150", "0x1.000001p0", "0x1.0000010000000001p0", "0x1.000003p0", "0xg", "0x.p1", "0x1p", "0x1p+",
             " 1", "\t1", "\n\r\v\f 2.5", " ", "", "-", "+", ".", "e5", "-.e5", "1 2", "1:2", "--1", "+-1", "1e5e5",
             "00000000000000000000000000000000000001", "0.00000000000000000000000000000000000000000000000001e50",
             "100000000000000000000000000000000000000000000000000e-50", "1e99999999999999999999", "1e-99999999999999999999",
             "0." + "0" * 400 + "1e401", "1" + "0" * 400 + "e-400", "9" * 200, "0." + "9" * 200]
    for _ in range(300):
        f = struct.unpack("<f", struct.pack("<I", rnd.getrandbits(32) & 0x7F7FFFFF))[0]
        cases.append(repr(float(f)))
        cases.append("%.9g" % f)
        cases.append("%.6g" % f)
        cases.append("%.4f" % (rnd.random() * 20))
    cases += ["0x1.400001p-148", "0x1.200001p-147", "0x1.400002p-148"]
    out = []
    for s in cases:
        v, n = O.strtof(s)
        bits = int(np.float32(v).view(np.uint32))
        entry = [s, bits, n]
        # glibc 2.39's HEX path drops the round bit of a 25-bit mantissa in the subnormal range
        # (e.g. 0x1.000001p-150 -> 0, IEEE nearest is 2^-149).  The IEEE result is what we pin;
        # the glibc value is kept beside it for the record.
        if s.strip().lstrip("+-").lower().startswith("0x") and n == len(s):
            ieee = hex_to_f32_bits(s)
            if ieee is not None and ieee != bits:
                entry = [s, ieee, n, {"glibc_2_39_bits": bits}]
        out.append(entry)
    return out


def hex_to_f32_bits(s):
    """exact IEEE round-to-nearest-even of a C hex-float string, by rational arithmetic"""
    import re
    from fractions import Fraction
    m = re.fullmatch(r"\s*([+-]?)0[xX]([0-9a-fA-F]*)\.?([0-9a-fA-F]*)(?:[pP]([+-]?\d+))?", s)
    if not m or not (m.group(2) or m.group(3)):
        return None
    sign, ip, fp, ex = m.groups()
    v = Fraction(int((ip or "0") + (fp or ""), 16), 16 ** len(fp or "")) * Fraction(2) ** int(ex or 0)
    sbit = 0x80000000 if sign == "-" else 0
    if v == 0:
        return sbit
    e = 0
    while v >= 2 ** (e + 1):
        e += 1
    while v < Fraction(2) ** e:
        e -= 1
    e = max(e, -126)
    q = v / Fraction(2) ** (e - 23)            # in units of the ulp at this exponent
    n = q.numerator // q.denominator
    r = q - n
    if r > Fraction(1, 2) or (r == Fraction(1, 2) and n % 2 == 1):
        n += 1
    if n >= 1 << 24:
        n >>= 1
        e += 1
    if e > 127:
        return sbit | 0x7f800000
    if n < 1 << 23:                              # subnormal (e == -126)
        return sbit | n
    return sbit | ((e + 127) << 23) | (n - (1 << 23))


def strtoll_vectors():
    cases = ["0", "-0", "+7", "123", "  42", "\t-9", "9223372036854775807", "9223372036854775808",
             "-9223372036854775808", "-9223372036854775809", "99999999999999999999999", "", "-", "+", "a", "12a",
             "0x10", "1e5", "007", " ", "--1"]
    return [[s, *O.strtoll(s)] for s in cases]


def main():
    with open(os.path.join(HERE, "parser_vectors.json"), "w") as f:
        json.dump(parser_vectors(), f, indent=1)
    with open(os.path.join(HERE, "scorer_case.json"), "w") as f:
        json.dump(scorer_case(), f, indent=1)
    with open(os.path.join(HERE, "strtof_vectors.json"), "w") as f:
        json.dump({"strtof": strtof_vectors(), "strtoll": strtoll_vectors()}, f)
    ref = "/root/reference/data"
    if os.path.isdir(ref):
        for src, dst in [("train_0", "criteo_excerpt.txt"), ("weight_0", "criteo_excerpt_weights.txt")]:
            with open(os.path.join(ref, src)) as fi, open(os.path.join(HERE, dst), "w") as fo:
                for i, line in enumerate(fi):
                    if i >= 2000:
                        break
                    fo.write(line)


if __name__ == "__main__":
    main()
