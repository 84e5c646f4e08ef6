"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/sav_abi.h declares (no compute calls without a GPU), the ctypes table matches the header,
the product path fails loudly without the library / a device, and the linen-compatible module
surface reproduces the reference's names, shapes and call signatures (shape propagation only, on
the 'meta' device -- the analogue of the reference's chex.assert_shape tests).
"""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "sav_abi.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sav_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def built_lib():
    import __graft_entry__ as G
    G.build()
    from sav import _abi
    return _abi


def test_header_declares_entry_points():
    names = _declared_functions()
    for must in ("sav_gemm", "sav_attn_fwd", "sav_attn_bwd", "sav_layernorm_fwd", "sav_layernorm_bwd",
                 "sav_patchify", "sav_embed_fwd", "sav_embed_bwd", "sav_softmax_xent", "sav_adam_step"):
        assert must in names


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib.LIB_PATH)
    for name in _declared_functions():
        assert hasattr(lib, name), f"{name} declared in include/sav_abi.h but not exported"
    assert built_lib.lib().sav_abi_version() == 3


def test_ctypes_table_matches_header(built_lib):
    assert sorted(built_lib.SIGNATURES) == _declared_functions()


def test_struct_layouts_agree_with_a_c_compiler(built_lib, tmp_path):
    """Compile include/sav_abi.h as plain C (it must be a C header) and compare sizeof/offsetof of
    the argument structs with the ctypes mirrors."""
    import subprocess
    src = tmp_path / "probe.c"
    src.write_text(f"""#include <stdio.h>
#include <stddef.h>
#include "{HEADER}"
int main(void) {{
  printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(sav_gemm_args), offsetof(sav_gemm_args, residual),
         sizeof(sav_attn_fwd_args), offsetof(sav_attn_fwd_args, scale),
         sizeof(sav_attn_bwd_args), offsetof(sav_attn_bwd_args, scale),
         sizeof(sav_bgemm_args), offsetof(sav_bgemm_args, a_mn),
         sizeof(sav_thsoftmax_args), offsetof(sav_thsoftmax_args, scale),
         sizeof(sav_conv2d_geom), offsetof(sav_conv2d_geom, ld));
  return 0;
}}""")
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    G, F, B = built_lib.GemmArgs, built_lib.AttnFwdArgs, built_lib.AttnBwdArgs
    BG, TS, CG = built_lib.BGemmArgs, built_lib.ThSoftmaxArgs, built_lib.Conv2dGeom
    assert got == [ctypes.sizeof(G), G.residual.offset, ctypes.sizeof(F), F.scale.offset,
                   ctypes.sizeof(B), B.scale.offset, ctypes.sizeof(BG), BG.a_mn.offset,
                   ctypes.sizeof(TS), TS.scale.offset, ctypes.sizeof(CG), CG.ld.offset]


def test_no_device_is_a_loud_error(built_lib):
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    assert built_lib.lib().sav_device_check() != 0
    assert b"CUDA" in built_lib.lib().sav_last_error() or b"device" in built_lib.lib().sav_last_error()


def test_missing_library_raises(monkeypatch, built_lib):
    monkeypatch.setattr(built_lib, "_lib", None)
    monkeypatch.setattr(built_lib, "LIB_PATH", "/nonexistent/libsav_b200.so")
    with pytest.raises(built_lib.SavError, match="no CPU or PyTorch fallback"):
        built_lib.lib()


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "self-attention-experiments-vision_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, os.path.join(dp, f)


# ------------------------------------------------------------------------- module surface (meta)
@pytest.mark.parametrize("name,img", [("vit_b_patch32", 224), ("vit_b_patch16", 224),
                                      ("deit_ti_patch16", 224), ("deit_s_patch16", 224)])
def test_logits_shape_like_reference_tests(name, img):
    """models/vit_test.py:13-26: init_with_output on ones[2,R,R,3] -> logits (2, 1000)."""
    from models import create_model
    model = create_model(name, num_classes=1000)
    x = torch.empty((2, img, img, 3), device="meta")
    logits, variables = model.init_with_output({"params": 0}, x, is_training=True)
    assert tuple(logits.shape) == (2, 1000)
    assert "params" in variables


def test_param_tree_matches_oracle_names_and_shapes():
    from models import create_model
    from oracle import vit_oracle as O
    model = create_model("deit_s_patch16")
    x = torch.empty((1, 224, 224, 3), device="meta")
    variables = model.init({"params": 0}, x, is_training=False)
    ours = {p: tuple(t.shape) for p, t in O.tree_leaves(variables["params"])}
    ref = {p: tuple(t.shape) for p, t in
           O.tree_leaves(O.synthetic_vit_params(O.vit_config("deit_s_patch16"), 224))}
    assert ours == ref


def test_module_fields_match_reference_signatures():
    """attention.py:11-18, ff.py:9-13, patch_embed.py:10-13, layerscale.py:14-15,
    stochastic_depth.py:7-8: dataclass field names/defaults are the public API."""
    import dataclasses
    from models.layers import (AttentionBlock, FFBlock, LayerScaleBlock, PatchEmbedBlock,
                               StochasticDepthBlock)
    f = {x.name: x.default for x in dataclasses.fields(AttentionBlock)}
    for k, v in dict(head_ch=None, out_ch=None, talking_heads=False, attn_dropout_rate=0.,
                     out_dropout_rate=0., use_bias=False).items():
        assert f[k] == v
    assert "num_heads" in f
    f = {x.name: x.default for x in dataclasses.fields(FFBlock)}
    assert f["expand_ratio"] is None and f["hidden_ch"] is None and f["dropout_rate"] == 0.
    f = {x.name: x.default for x in dataclasses.fields(PatchEmbedBlock)}
    assert f["use_bias"] is False
    assert {x.name for x in dataclasses.fields(LayerScaleBlock)} >= {"eps"}
    f = {x.name: x.default for x in dataclasses.fields(StochasticDepthBlock)}
    assert f["drop_rate"] == 0. and f["scale_by_keep"] is True


def test_ffblock_requires_ratio_or_hidden():
    from models.layers import FFBlock
    from sav.tape import Var
    x = Var(torch.empty((2, 5, 32), dtype=torch.bfloat16, device="meta"))
    with pytest.raises(ValueError):
        FFBlock().init({"params": 0}, x, is_training=False, device="meta")


def test_unknown_model_name():
    """create_model.py:214-215: RuntimeError('Model not found.'); every name the reference knows builds."""
    from models import create_model
    with pytest.raises(RuntimeError, match="Model not found"):
        create_model("resnet50")
    ref_names = ['vit_b_patch32', 'vit_b_patch16', 'vit_l_patch32', 'vit_l_patch16', 'botnet_t3', 'botnet_t4',
                 'botnet_t5', 'tnt_s_patch16', 'tnt_b_patch16', 'ceit_t', 'ceit_s', 'ceit_b', 'cait_xxs_24',
                 'cait_xxs_36', 'cait_xs_24', 'cait_xs_36', 'cait_s_24', 'cait_s_36', 'cait_s_48', 'cait_m_24',
                 'cait_m_36', 'cait_m_48', 'cvt-13', 'cvt-21', 'cvt-w24', 'mixer_s_patch32', 'mixer_s_patch16', 'mixer_b_patch32',
                 'mixer_l_patch32', 'mixer_l_patch16']
    for name in ref_names:
        assert create_model(name, num_classes=10) is not None, name
