"""CLIP text encoder on the b200ad kernels (-m gpu) against transformers' own CLIPTextModel -- the
reference's dependency for this step (/root/reference/requirements.txt:7), importable on this image -- and against the
restated oracle, on identical bf16-representable random-init weights and token ids."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


@pytest.mark.parametrize("mode,tol", [("bf16", 1e-2), ("fp32", 1e-4)])
def test_clip_text_encoder_matches_transformers(mode, tol):
    transformers = pytest.importorskip("transformers")
    from b200ad.clip import CLIPTextEncoder
    from oracle.clip_text import make_oracle_clip
    oracle = make_oracle_clip(seed=2)                            # full 12-layer ViT-L/14 text tower, weights rounded to bf16
    tc = transformers.CLIPTextConfig(hidden_size=768, intermediate_size=3072, num_hidden_layers=12,
                                     num_attention_heads=12, max_position_embeddings=77, vocab_size=49408,
                                     hidden_act="quick_gelu")
    lib = transformers.CLIPTextModel(tc).eval()
    lib.load_state_dict(oracle.state_dict(), strict=False)
    prod = CLIPTextEncoder(mode=mode)
    prod.load_state_dict(oracle.state_dict())
    prod.cuda()
    ids = torch.randint(0, 49408, (2, 77), generator=torch.Generator().manual_seed(0))
    ids[1, 20:] = 49407                                          # a padded prompt, as the tokenizer produces
    with torch.no_grad():
        ref = lib(ids)[0]
    out = prod(ids.cuda())
    assert out.shape == (2, 77, 768) and out.dtype == torch.float32
    e_lib, e_orc = rel(out.cpu(), ref), rel(out.cpu(), oracle(ids))
    print(f"clip text encoder {mode}: vs transformers {e_lib:.3e}, vs oracle {e_orc:.3e}")
    assert e_lib < tol and e_orc < tol
    # shorter sequences and batch 1 (prompt_embeds for one prompt)
    assert rel(prod(ids[:1, :32].cuda()).cpu(), oracle(ids[:1, :32])) < tol


def test_pipeline_uses_local_text_encoder_when_present(tmp_path):
    """encode_prompt runs the b200ad CLIP encoder when the checkpoint directory provides text_encoder/ (+ tokenizer/);
    here the encoder weights are saved locally and the tokenizer is replaced by a stub with CLIPTokenizer's call shape."""
    from safetensors.torch import save_file
    from b200ad.clip import CLIPTextEncoder
    from b200ad.pipeline import _LocalClip
    from oracle.clip_text import ClipTextConfig, make_oracle_clip
    import json
    cfg = ClipTextConfig(num_hidden_layers=2)
    oracle = make_oracle_clip(seed=3, cfg=cfg)
    d = tmp_path / "text_encoder"
    d.mkdir()
    save_file({k: v.contiguous() for k, v in oracle.state_dict().items()}, str(d / "model.safetensors"))
    (d / "config.json").write_text(json.dumps({"num_hidden_layers": 2, "hidden_act": "quick_gelu", "hidden_size": 768,
                                               "intermediate_size": 3072, "num_attention_heads": 12,
                                               "max_position_embeddings": 77, "vocab_size": 49408}))
    enc = CLIPTextEncoder.from_local(str(d))

    class Tok:
        model_max_length = 77

        def __call__(self, prompt, **kw):
            g = torch.Generator().manual_seed(len(prompt))
            return type("E", (), {"input_ids": torch.randint(0, 49408, (1, 77), generator=g)})()
    clip = _LocalClip(Tok(), enc)
    emb = clip("a corgi walking on the beach")
    assert emb.shape == (1, 77, 768) and emb.is_cuda
    ids = Tok()("a corgi walking on the beach").input_ids
    assert rel(emb.cpu(), oracle(ids)) < 1e-2
