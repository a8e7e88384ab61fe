"""Generates the committed golden fixtures from the upstream artefacts.  Run in the build container only
(/root/reference does not exist on the GPU box):

    python tests/golden/make_golden.py

Outputs (small, committed):
  outbmp_block_means.npz   8x8-block means of /root/reference/out.bmp (the only upstream render: 800x800, 1024 spp of
                           test_scene.txt, src/main.rs:13-60) decoded to LINEAR light with the reference's own sRGB table
                           (src/color.rs:75-332), plus the fraction of saturated (code 255) channel values per block and
                           the 122-byte header.  100 x 100 x 3 float32.
  srgb_tables.npz          SRGB_VALUES[256] / SRGB_AVERAGE[255] parsed from src/color.rs:75-332,335-591 -- the literals the
                           quantiser is defined by; the oracle's regenerated tables are checked against them bit for bit.
  test_scene_sha256.txt    digest of upstream test_scene.txt, so tests can tell that REF_TEXT_SCENE restates the same scene.
"""
import hashlib
import os
import re
import sys

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def parse_table(src, name):
    m = re.search(r"%s\s*:\s*\[f64;\s*(\d+)\]\s*=\s*\[(.*?)\];" % name, src, re.S)
    n = int(m.group(1))
    vals = [float(x) for x in re.findall(r"[-+]?\d+\.?\d*(?:[eE][-+]?\d+)?", m.group(2))]
    assert len(vals) == n, (name, len(vals), n)
    return np.array(vals, dtype=np.float64)


def main():
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference")
    color_rs = open(os.path.join(REF, "src", "color.rs")).read()
    values = parse_table(color_rs, "SRGB_VALUES")
    average = parse_table(color_rs, "SRGB_AVERAGE")
    np.savez_compressed(os.path.join(HERE, "srgb_tables.npz"), values=values, average=average)

    raw = open(os.path.join(REF, "out.bmp"), "rb").read()
    header = np.frombuffer(raw[:122], dtype=np.uint8)
    w = int.from_bytes(raw[18:22], "little")
    h = int.from_bytes(raw[22:26], "little")
    off = int.from_bytes(raw[10:14], "little")
    bw = (3 * w + 3) & ~3
    rows = np.frombuffer(raw[off:off + bw * h], dtype=np.uint8).reshape(h, bw)[:, :3 * w].reshape(h, w, 3)  # row 0 = bottom, BGR
    lin = values[rows]                                                       # linear light, BGR order
    blocks = lin.reshape(h // 8, 8, w // 8, 8, 3).mean(axis=(1, 3)).astype(np.float32)
    sat = (rows == 255).reshape(h // 8, 8, w // 8, 8, 3).mean(axis=(1, 3)).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "outbmp_block_means.npz"), header=header, width=w, height=h, block=8, linear_bgr=blocks,
                        saturated=sat, mean_code_bgr=rows.reshape(-1, 3).mean(axis=0))
    with open(os.path.join(HERE, "test_scene_sha256.txt"), "w") as f:
        f.write(hashlib.sha256(open(os.path.join(REF, "test_scene.txt"), "rb").read()).hexdigest() + "\n")
    print("wrote fixtures:", w, h, blocks.shape, rows.reshape(-1, 3).mean(axis=0))


if __name__ == "__main__":
    main()
