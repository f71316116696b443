"""developer: short-row (Hkv_local <= 2) KV load hint sweep for the decode attention kernel (knob attn_ld_hint)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import kbench
from iaas_sglang_b200 import _lib
for hint in (0, 1):
    _lib.tune_set("attn_ld_hint", hint)
    print(f"### attn_ld_hint={hint}", flush=True)
    kbench.attn_case("cfg2tp8", 512, 4, 1, 128, 4096)
    kbench.attn_case("cfg4tp8", 256, 8, 1, 128, 8192, shared=2048)
    kbench.attn_case("cfg3tp4", 128, 10, 2, 128, 4096)
_lib.tune_set("attn_ld_hint", 0)
