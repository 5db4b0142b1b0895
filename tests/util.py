"""Shared helpers for the test-suite."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def product():
    """The product package (its directory name has a hyphen)."""
    return importlib.import_module("image-webp_b200")


def vp8_payload(data):
    """Payload of the first 'VP8 ' chunk of a WebP file."""
    b = bytes(data)
    i = b.find(b"VP8 ")
    n = int.from_bytes(b[i + 4:i + 8], "little")
    return b[i + 8:i + 8 + n]


def chunk_payload(data, fourcc):
    b = bytes(data)
    i = b.find(fourcc)
    n = int.from_bytes(b[i + 4:i + 8], "little")
    return b[i + 8:i + 8 + n]


LOSSY_STILL = (["gallery1/%d" % i for i in range(1, 6)] + ["gallery2/%d_webp_a" % i for i in range(1, 6)]
               + ["regression/dark"])

# feature-coverage cases for the synthetic writer (harness/synth.vp8_frame keyword sets)
SYNTH_CASES = {
    "simple_150x90": dict(width=150, height=90, seed=1, filter_type=1, filter_level=20),
    "normal_sharp3_2parts_131x77": dict(width=131, height=77, seed=2, filter_type=0, filter_level=30, sharpness=3, log2_parts=1),
    "abs_segments_level63_sharp7_4parts_33x17": dict(width=33, height=17, seed=3, filter_type=0, filter_level=63, sharpness=7,
                                                     log2_parts=2, segments=1, seg_quant=(20, 35, 50, 70), seg_lf=(10, 20, 30, 45),
                                                     seg_probs=(128, 100, 200)),
    "delta_segments_lfdeltas_8parts_skip_160x160": dict(width=160, height=160, seed=4, filter_type=0, filter_level=12, segments=2,
                                                        seg_quant=(-10, 5, 20, -3), seg_lf=(3, -5, 10, 0), seg_probs=(90, 128, -1),
                                                        lf_adj=2, ref_delta=(2, 0, 0, 0), mode_delta=(-3, 0, 0, 0), log2_parts=3,
                                                        prob_skip=180, p_skip=0.2),
    "extreme_levels_q127_64x64": dict(width=64, height=64, seed=5, profile="extreme", yac_abs=127, filter_type=0, filter_level=40, sharpness=5),
    "all_tm_64x48": dict(width=64, height=48, seed=6, profile="all_tm", filter_level=0),
    "bpred_ld_vl_80x80": dict(width=80, height=80, seed=7, profile="bpred_ld_vl", filter_type=1, filter_level=50, sharpness=6),
    "all_skipped_48x48": dict(width=48, height=48, seed=8, profile="zero_tails", filter_type=0, filter_level=25, prob_skip=128, p_skip=1.0),
    "zero_tails_48x48": dict(width=48, height=48, seed=9, profile="zero_tails", filter_type=0, filter_level=25, prob_skip=128, p_skip=0.3),
    "one_pixel": dict(width=1, height=1, seed=10),
    "uniform_17x33": dict(width=17, height=33, seed=11, profile="uniform", filter_type=0, filter_level=33),
    "prob_updates_seg_nomap_96x64": dict(width=96, height=64, seed=12, profile="uniform", update_probs=1, filter_type=0, filter_level=20,
                                         segments=3, seg_quant=(60, 0, 0, 0), seg_lf=(40, 0, 0, 0)),
    "seg_map_nodata_lfadj_noupdate_96x64": dict(width=96, height=64, seed=13, segments=4, seg_probs=(100, 100, 100), filter_type=0,
                                                filter_level=17, lf_adj=1),
    "cfg2_like_512x512": dict(width=512, height=512, seed=14, filter_type=1, filter_level=25, prob_skip=200, p_skip=0.05),
    "wide_row_640x16": dict(width=640, height=16, seed=15, filter_type=0, filter_level=22),
    "tall_column_16x640": dict(width=16, height=640, seed=16, filter_type=1, filter_level=22),
    "level0_segments_lf_50x50": dict(width=50, height=50, seed=17, filter_type=0, filter_level=0, segments=1, seg_quant=(20, 30, 40, 50),
                                     seg_lf=(20, 30, 40, 50), seg_probs=(128, 128, 128)),
    "even_odd_254x255": dict(width=254, height=255, seed=18, filter_type=0, filter_level=35, sharpness=1, profile="uniform"),
}
