#!/usr/bin/env python
"""Shrinks the raw outputs of make_golden.py into the committed fixtures: all block keys, visible
keys, visibility flags and traced images, a 64-bit checksum of every block's 2560 voxel words,
and the full voxel words of every 8th block (sorted by key).

    python tests/golden/pack_golden.py gpurun_out/golden tests/golden
"""
import os
import sys

import numpy as np


def block_checksums(voxels):
    """(n, 512, 5) int32 -> (n,) uint64: position-weighted sum of the words, wraps mod 2^64."""
    w = voxels.reshape(len(voxels), -1).astype(np.uint32).astype(np.uint64)
    k = (np.arange(w.shape[1], dtype=np.uint64) * np.uint64(2654435761) + np.uint64(12345)) | np.uint64(1)
    with np.errstate(over="ignore"):
        return (w * k).sum(axis=1, dtype=np.uint64)


def main(src, dst):
    for f in sorted(os.listdir(src)):
        if not f.endswith(".npz"):
            continue
        d = np.load(os.path.join(src, f))
        vox = d["voxels"]
        np.savez_compressed(os.path.join(dst, f), keys=d["keys"], visible_keys=d["visible_keys"],
                            vis_flags=d["vis_flags"], checksums=block_checksums(vox), sample_voxels=vox[::8],
                            depth=d["depth"], color=d["color"], normal=d["normal"])
        print(f, os.path.getsize(os.path.join(dst, f)))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
