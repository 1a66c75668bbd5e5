"""Run-time switches that are not part of the reference's signatures.

numpy_semantics: which NumPy's np.percentile arithmetic the thresholds reproduce.
  'legacy_1_22' -- the reference's pin (requirements.txt:2, numpy==1.22.4); float64 index/lerp,
                   float32 compare.  Default, because it is what produced the reference's logs.
  'nep50_2x'    -- NumPy >= 2 (the NumPy installed on the GPU box); float32 index/lerp.
Override with the environment variable WT_NUMPY_SEMANTICS or by assigning to this attribute.

async_save: save_model snapshots the state dict and writes it from a background thread, so the next
pruning method starts while safetensors serialises (once the kernels take microseconds, `save_pretrained` IS the
critical path of main_pruning.py:175-209).  Off by default: the reference returns with the files on disk.
utils.wait_for_saves() joins; it also runs at interpreter exit.

arithmetic: 'faithful' (default: fp32 multiply then add, no FMA -- the arithmetic of the oracle and of every
  bit-exactness claim) or 'fma_taporder' (named fast mode: the same taps in the same order, every multiply-add fused; for
  the fp32-issue-bound decompositions; masks / zero counts may differ from the reference's).  WT_ARITHMETIC overrides.

verbose: re-enables the reference's per-layer debug prints (dwt_pruning.py:29-30,121-122).
"""
import os

numpy_semantics = os.environ.get('WT_NUMPY_SEMANTICS', 'legacy_1_22')
verbose = os.environ.get('WT_VERBOSE', '0') not in ('0', '', 'false', 'False')
async_save = os.environ.get('WT_ASYNC_SAVE', '0') not in ('0', '', 'false', 'False')
arithmetic = os.environ.get('WT_ARITHMETIC', 'faithful')
