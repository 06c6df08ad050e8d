"""Regenerates tests/golden/*.npz from the oracle (seeded, small).  Run from the repo root:
    python tests/golden/make_golden.py
These vectors come from the ORACLE (they keep its behaviour fixed over time and give the GPU tests fixed inputs and
outputs for the rows whose reference arithmetic sits in absent third-party packages); the fixtures written by the
reference's own code are ref_*.npz, made by make_reference_golden.py; hand-derived known answers live in
tests/test_oracle_golden.py."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from oracle import paths as P  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    rng = np.random.default_rng(20261106)
    # 1. pre-processing + tiling, both conventions
    img = O.synthetic_image((1, 1, 96, 120), 10, seed=11)
    np.savez_compressed(
        os.path.join(OUT, "pre_tiles.npz"), img=img,
        scale_range=O.scale_range(img, 1.0, 99.8, axes=(1, 2, 3)),
        zmuv=O.zero_mean_unit_variance(img, axes=(1, 2, 3)),
        fibsem_norm=O.fibsem_normalize(img[0, 0]),
        quantiles=np.nanquantile(img, np.array([0.01, 0.5, 0.998])),
        crop_tiles=P.crop_pre(img, (48, 48), (8, 8)),
        fibsem_tiles=P.c2_pre(img, 48, 16),
    )
    # 2. stitching + post-processing on fixed pseudo-logits
    tiles = P.c2_pre(img, 48, 16)
    logits = (np.sin(tiles * 7.0) * 3.0 + rng.standard_normal(tiles.shape).astype(np.float32) * 0.1).astype(np.float32)
    prob, mask, labels = P.c2_post(logits, (96, 120), 48, 16, 2)
    ctiles = P.crop_pre(img, (48, 48), (8, 8))
    clog = (np.cos(ctiles * 5.0) * 2.0).astype(np.float32)
    cprob, cmask, clabels = P.crop_post(clog, (1, 1, 96, 120), (48, 48), (8, 8))
    np.savez_compressed(os.path.join(OUT, "stitch_post.npz"), blend_logits=logits, blend_prob=prob, blend_mask=mask,
                        blend_labels=labels, crop_logits=clog, crop_prob=cprob, crop_mask=cmask, crop_labels=clabels)
    # 3. labels in 2-D and 3-D for every connectivity
    m2 = rng.random((40, 70)) < 0.45
    m3 = rng.random((9, 12, 70)) < 0.3
    np.savez_compressed(os.path.join(OUT, "labels.npz"), m2=m2, m3=m3, l2_c1=O.label(m2, 1), l2_c2=O.label(m2, 2),
                        l3_c1=O.label(m3, 1), l3_c2=O.label(m3, 2), l3_c3=O.label(m3, 3))
    # 4. metric
    a = O.gaussian_blobs((80, 100), 14, seed=3)
    t = O.label(a > 0.5, 2)
    p = O.label(a + 0.25 * rng.standard_normal(a.shape) > 0.5, 2)
    ap, tp, fp, fn = O.average_precision([t], [p], threshold=[0.5, 0.75, 0.9])
    np.savez_compressed(os.path.join(OUT, "metric.npz"), true=t, pred=p, overlap=O.label_overlap(t, p), ap=ap, tp=tp,
                        fp=fp, fn=fn)
    # 5. fibsem instance post-processing (size filter → closing → EDT → peaks → markers → flood): touching blobs, and a
    #    mask with twin discs and straight bars whose seeds tie in height (the two tie orders differ there)
    blobs = np.clip(O.gaussian_blobs((160, 200), 14, sigma=(6.0, 14.0), seed=5) * 1.3, 0, 1).astype(np.float32)
    yy, xx = np.mgrid[:120, :160]
    twins = (((yy - 40) ** 2 + (xx - 40) ** 2 <= 15**2) | ((yy - 40) ** 2 + (xx - 66) ** 2 <= 15**2)).astype(np.float32)
    twins[80:92, 20:140] = 1.0
    out = {}
    for name, prob in (("blobs", blobs), ("twins", twins)):
        parts = {}
        fifo = O.prob_to_instances(prob, ties="fifo", parts=parts)
        out.update({f"{name}_prob": prob, f"{name}_closed": parts["closed"], f"{name}_dist2": np.rint(parts["dist"] ** 2).astype(np.int32),
                    f"{name}_peaks": parts["coords"].astype(np.int32), f"{name}_markers": parts["markers"],
                    f"{name}_fifo": fifo, f"{name}_heap": O.prob_to_instances(prob, ties="heap")})
    np.savez_compressed(os.path.join(OUT, "instances.npz"), **out)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
