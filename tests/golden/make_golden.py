"""Generates the committed golden vectors under tests/golden/.

Run in the authoring container, where /root/reference (the reference checkout) and cv2 exist:
    python tests/golden/make_golden.py
The GPU box has neither /root/reference nor a need for cv2: tests only read the .npz files written here.

Sources of truth
  * cv2 4.x  cvtColor(BGR2HSV) + inRange + morphologyEx -- the library calls behind hsvThresholdingDetector
    (pcs_detection/include/pcs_detection/hsv_thresholding.h:55-59) and the histogram detector's opening
    (pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:40-41)
  * the reference's test fixtures pcs_scan_integration/test/{box_2m.bt, plane_4m.stl, box_2m.ply}, stored here in
    decoded form (leaf keys / welded indexed mesh) plus a SHA-256 of the original bytes.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
REF = "/root/reference/pcs_scan_integration/test"




def hsv_golden():
    import cv2

    lower, upper = (95, 95, 40), (115, 220, 150)
    # all 2^24 colours as a 4096 x 4096 image: r slowest, then g, b fastest
    r = np.arange(256, dtype=np.uint8)
    img = np.stack([np.broadcast_to(r[None, None, :], (256, 256, 256)),
                    np.broadcast_to(r[None, :, None], (256, 256, 256)),
                    np.broadcast_to(r[:, None, None], (256, 256, 256))], axis=-1).reshape(4096, 4096, 3).copy()
    hsv = cv2.cvtColor(img, cv2.COLOR_BGR2HSV)
    mask = cv2.inRange(hsv, lower, upper)
    out = {
        "cv2_version": np.array(cv2.__version__),
        "exhaustive_mask_sha256": np.array(hashlib.sha256(mask.tobytes()).hexdigest()),
        "exhaustive_hsv_sha256": np.array(hashlib.sha256(hsv.tobytes()).hexdigest()),
        "exhaustive_positive": np.array(int((mask == 255).sum())),
        "exhaustive_mask_packed": np.packbits(mask.reshape(-1) != 0),  # 2 MiB before compression, ~20 KiB after
    }
    rng = np.random.default_rng(20191002)
    shapes = [(37, 53), (8, 128), (9, 129), (64, 200), (1, 17), (240, 320)]
    for i, (h, w) in enumerate(shapes):
        im = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        # paint blobs of an in-range colour so that the opening has something to keep
        for _ in range(6):
            y0, x0 = rng.integers(0, h), rng.integers(0, w)
            hh, ww = rng.integers(1, 12), rng.integers(1, 12)
            im[y0:y0 + hh, x0:x0 + ww] = (140, 90, 50)
        m = cv2.inRange(cv2.cvtColor(im, cv2.COLOR_BGR2HSV), lower, upper)
        mo = cv2.morphologyEx(m, cv2.MORPH_OPEN, np.ones((5, 5), np.uint8))
        out[f"img{i}"] = im
        out[f"mask{i}"] = m
        out[f"open{i}"] = mo
    out["n_images"] = np.array(len(shapes))
    np.savez_compressed(os.path.join(HERE, "hsv_cv2.npz"), **out)
    print("hsv_cv2.npz: positives", int(out["exhaustive_positive"]))


def fixture_golden():
    from oracle import oracle as O

    bt = open(os.path.join(REF, "box_2m.bt"), "rb").read()
    tree = O.Octree.read_bt(bt)
    kx, ky, kz, d, lo = tree.leaves()
    stl = open(os.path.join(REF, "plane_4m.stl"), "rb").read()
    v, f = O.load_stl_binary(stl)
    ply = open(os.path.join(REF, "box_2m.ply"), "rb").read()
    pv, pf = O.load_ply_ascii(ply)
    np.savez_compressed(
        os.path.join(HERE, "reference_fixtures.npz"),
        bt_sha256=np.array(hashlib.sha256(bt).hexdigest()), bt_size=np.array(len(bt)),
        bt_resolution=np.array(tree.resolution), bt_num_nodes=np.array(tree.num_nodes()),
        bt_kx=kx, bt_ky=ky, bt_kz=kz, bt_depth=d, bt_logodds=lo,
        stl_sha256=np.array(hashlib.sha256(stl).hexdigest()),
        plane_verts=v.astype(np.float32), plane_faces=f[:, 1:].astype(np.int32),
        ply_sha256=np.array(hashlib.sha256(ply).hexdigest()),
        box_verts=pv.astype(np.float32), box_faces=pf[:, 1:].astype(np.int32),
        # octomap_mesh_masking_unit.cpp:76-77,85-86,158-159,170-171,182-183
        golden_plane_nv=np.array(6273), golden_plane_nt=np.array(12288), golden_box_nv=np.array(8),
        golden_box_nt=np.array(12), golden_outside_tris=np.array(11400), golden_inside_entries_bullet=np.array(1975),
        golden_colorized=np.array([6273, 12288]),
    )
    print("reference_fixtures.npz: plane", v.shape, f.shape, "box", pv.shape, pf.shape, "bt leaves", len(kx))


def backproject_golden():
    """cv2 outputs of HistogramBackprojection.annotate_image
    (pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:22-52) for trained histograms of
    several bin counts."""
    import cv2

    def annotate(img, hist):
        hsv = cv2.cvtColor(img, cv2.COLOR_BGR2HSV)
        dst = cv2.calcBackProject([hsv], [0, 1], hist, [0, 180, 0, 255], 1)
        disc = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (5, 5))
        cv2.filter2D(dst, -1, disc, dst)
        _, t = cv2.threshold(dst, 30, 150, 0)
        c = cv2.morphologyEx(t, cv2.MORPH_OPEN, np.ones((5, 5), np.uint8))
        cv2.normalize(c, c, 0, 1, cv2.NORM_MINMAX, cv2.CV_8UC3)
        c = 255 * c
        return cv2.merge((c, c, c))

    rng = np.random.default_rng(20191002)
    out = {}
    cases = []
    for hi, (nh, ns) in enumerate([(180, 256), (30, 32), (50, 60)]):
        sample = np.zeros((40, 40, 3), np.uint8)
        sample[:] = (140, 90, 50)
        sample = (sample.astype(int) + rng.integers(-20, 20, sample.shape)).clip(0, 255).astype(np.uint8)
        hist = cv2.calcHist([cv2.cvtColor(sample, cv2.COLOR_BGR2HSV)], [0, 1], None, [nh, ns], [0, 180, 0, 255])
        cv2.normalize(hist, hist, 0, 255, cv2.NORM_MINMAX)
        out[f"hist{hi}"] = hist
        for si, shape in enumerate([(37, 53), (3, 3), (120, 160)]):
            img = rng.integers(0, 256, (*shape, 3), dtype=np.uint8)
            for _ in range(5):
                y0, x0 = rng.integers(0, shape[0]), rng.integers(0, shape[1])
                img[y0:y0 + rng.integers(1, 30), x0:x0 + rng.integers(1, 30)] = (140, 90, 50)
            img[..., 1][img[..., 1] == 90] = 255
            out[f"img{hi}_{si}"] = img
            out[f"out{hi}_{si}"] = annotate(img, hist)
            cases.append((hi, si))
        const = np.zeros((20, 20, 3), np.uint8)
        const[:] = (140, 90, 50)
        out[f"img{hi}_c"] = const
        out[f"out{hi}_c"] = annotate(const, hist)  # all detected -> min == max -> normalises to zeros
    out["cases"] = np.array(cases)
    np.savez_compressed(os.path.join(HERE, "backproject_cv2.npz"), **out)
    print("backproject_cv2.npz:", len(cases), "cases")


if __name__ == "__main__":
    backproject_golden()
    hsv_golden()
    fixture_golden()
