"""Small shapes through every kernel family, for compute-sanitizer (tools/sanitize.sh): dense causal, GQA pack, varlen,
paged dual-stream decode + combine, ragged decode (schedule kernel), the 2-CTA variant, the four-role variant, FP8,
append-KV + rotary, the no-alloc entry point.  Results are only checked for finiteness here (parity is pytest's job)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch  # noqa: E402
import flash_attn_interface as F  # noqa: E402

dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
mk = lambda *s, dt=torch.bfloat16: torch.randn(*s, generator=g, device=dev, dtype=torch.bfloat16).to(dt)  # noqa: E731


def done(name, out):
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all(), name
    print("ok", name, flush=True)


done("dense causal", F.flash_attn_func(mk(1, 300, 4, 128), mk(1, 300, 2, 128), mk(1, 300, 2, 128), causal=True))
done("pack_gqa splits", F.flash_attn_func(mk(1, 40, 8, 128), mk(1, 700, 2, 128), mk(1, 700, 2, 128), pack_gqa=True, num_splits=3))
done("hdim 64 local", F.flash_attn_func(mk(1, 260, 2, 64), mk(1, 260, 2, 64), mk(1, 260, 2, 64), window_size=(50, 10)))
done("hdim 256", F.flash_attn_func(mk(1, 200, 2, 256), mk(1, 200, 1, 256), mk(1, 200, 1, 256), causal=True))
cu = torch.tensor([0, 100, 357], dtype=torch.int32, device=dev)
done("varlen", F.flash_attn_varlen_func(mk(357, 4, 128), mk(357, 2, 128), mk(357, 2, 128), cu, cu, 257, 257, causal=True))
kp, vp = mk(8, 128, 2, 128), mk(8, 128, 2, 128)
table = torch.randperm(8, device=dev).to(torch.int32).view(2, 4)
lens = torch.tensor([512, 130], dtype=torch.int32, device=dev)
done("paged ragged decode", F.flash_attn_with_kvcache(mk(2, 1, 8, 128), kp, vp, cache_seqlens=lens, page_table=table))
kc, vc = mk(2, 256, 2, 128), mk(2, 256, 2, 128)
cos, sin = mk(300, 64), mk(300, 64)
done("append + rotary", F.flash_attn_with_kvcache(mk(2, 3, 4, 128), kc, vc, k=mk(2, 3, 2, 128), v=mk(2, 3, 2, 128), rotary_cos=cos, rotary_sin=sin,
                                                   cache_seqlens=torch.tensor([100, 7], dtype=torch.int32, device=dev), causal=True))
e4 = torch.float8_e4m3fn
done("fp8", F.flash_attn_func(mk(1, 300, 4, 128, dt=e4), mk(1, 300, 2, 128, dt=e4), mk(1, 300, 2, 128, dt=e4), causal=True))
os.environ["FA100_PAIR"] = "1"
done("2-CTA variant", F.flash_attn_func(mk(1, 600, 2, 128), mk(1, 600, 2, 128), mk(1, 600, 2, 128), causal=True))
os.environ["FA100_PAIR"] = "0"
os.environ["FA100_FOUR_ROLE"] = "1"
done("four-role variant", F.flash_attn_func(mk(1, 300, 4, 128), mk(1, 300, 2, 128), mk(1, 300, 2, 128), causal=True))
del os.environ["FA100_FOUR_ROLE"]
print("all cases ran", flush=True)
