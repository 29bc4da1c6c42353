"""Synthetic streams shared by the CPU (oracle) and GPU (CUDA) parity tests: small versions of BASELINE configs 2-5
plus bit-depth / channel-count / edge cases, all produced by tools/flacgen (deterministic seeds)."""
from tools import flacgen as fg

CASES = {
    "c2_stereo16_lpc8_midside": lambda: fg.c2(12.0),
    "c3_8ch24_lpc32_po7": lambda: fg.c3(1.5),
    "c4_variable_blocksize": lambda: fg.c4_variable(8.0),
    "c4_32bit_rice2_escapes": lambda: fg.c4_32bit(3.0),
    "c4_wasted_bits_mixed": lambda: fg.c4_wasted(8.0),
    "c4_blocksize_65535": lambda: fg.c4_bs65535(6.0),
    "c5_file_0": lambda: fg.c5_file(0),
    "c5_file_1": lambda: fg.c5_file(1),
    "c5_file_2": lambda: fg.c5_file(2),
    "c5_file_3": lambda: fg.c5_file(3),
    "bps8_fixed_lowprec": lambda: fg.params(bps=8, lpc_order=2, qlp_precision=6, stereo_mode=4, subframe_mix=1,
                                           n_samples=100000, amplitude=0.3, noise=0.01, seed=11),
    "bps12_midside": lambda: fg.params(bps=12, lpc_order=6, qlp_precision=10, stereo_mode=3, n_samples=100000,
                                       amplitude=0.3, noise=0.004, seed=12),
    "bps20_mono_rice2": lambda: fg.params(bps=20, channels=1, lpc_order=10, qlp_precision=14, rice_method=1,
                                          n_samples=100000, seed=13),
    "bps24_stereo_leftside": lambda: fg.params(bps=24, lpc_order=12, qlp_precision=15, rice_method=1, stereo_mode=1,
                                               sample_rate=96000, n_samples=120000, noise=0.0003, seed=14),
    "bps15_sideright": lambda: fg.params(bps=15, lpc_order=4, qlp_precision=9, stereo_mode=2, n_samples=60000, seed=15),
    "unknown_total": lambda: fg.params(set_total=0, stereo_mode=4, n_samples=50000, seed=16),
    "ch3_partition8": lambda: fg.params(channels=3, min_partition_order=8, max_partition_order=8, n_samples=90000, seed=17),
    "ch6_mixed": lambda: fg.params(channels=6, subframe_mix=1, sample_rate=48000, n_samples=48000, seed=18),
    "tiny_blocks_16": lambda: fg.params(block_size=16, lpc_order=2, qlp_precision=7, max_partition_order=2,
                                        stereo_mode=4, n_samples=16 * 700 + 5, seed=19),
    "one_sample_last_frame": lambda: fg.params(block_size=1024, n_samples=1024 * 9 + 1, stereo_mode=3, seed=20),
    "escapes_everywhere": lambda: fg.params(escape_every=2, max_partition_order=6, stereo_mode=3, n_samples=80000, seed=21),
    "order32_16bit": lambda: fg.params(lpc_order=32, qlp_precision=10, block_size=4608, max_partition_order=6,
                                       n_samples=60000, seed=22),
}


def make(name):
    return fg.generate(CASES[name](), want_pcm=True)
