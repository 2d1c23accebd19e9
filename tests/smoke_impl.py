"""Body of ``__graft_entry__.smoke()``: one decode step of the hot-path ops on cuda:0, checked
against the CPU oracle (the oracle is the checker here, never the thing shipped)."""
import torch

from oracle import ops as O

BF16 = torch.bfloat16


def run(dev):
    import nano_qwen3tts_vllm_b200._native as nat
    g = torch.Generator().manual_seed(0)
    b, h, kvh, hd, page = 4, 16, 8, 128, 256
    lens = [70, 300, 1, 513]
    qkv = torch.randn(b, (h + 2 * kvh) * hd, generator=g).to(BF16)
    qw = (1 + 0.05 * torch.randn(hd, generator=g)).to(BF16)
    kw = (1 + 0.05 * torch.randn(hd, generator=g)).to(BF16)
    kc = torch.randn(8, page, kvh, hd, generator=g).to(BF16)
    vc = torch.randn(8, page, kvh, hd, generator=g).to(BF16)
    bt = torch.tensor([[5, -1, -1], [2, 7, -1], [0, -1, -1], [1, 3, 4]], dtype=torch.int32)
    cl = torch.tensor(lens, dtype=torch.int32)
    slots = torch.tensor([bt[i, (n - 1) // page] * page + (n - 1) % page for i, n in enumerate(lens)], dtype=torch.int32)
    pos = torch.tensor(lens, dtype=torch.int64)
    inv = O.make_inv_freq(hd, 1e6)
    q, k, v = qkv[:, :2048].view(b, h, hd), qkv[:, 2048:3072].view(b, kvh, hd), qkv[:, 3072:].view(b, kvh, hd)
    # oracle
    q_ref, k_ref = O.qknorm_rope(q, k, qw, kw, 1e-6, pos, inv)
    kc_ref, vc_ref = kc.clone(), vc.clone()
    O.store_kvcache(k_ref, v, kc_ref, vc_ref, slots)
    o_ref = O.paged_attn_decode(q_ref, kc_ref, vc_ref, cl, bt, hd ** -0.5)
    # device
    d = qkv.to(dev)
    kcd, vcd = kc.to(dev), vc.to(dev)
    qd, _ = nat.qknorm_rope_kvwrite(d[:, :2048].view(b, h, hd), d[:, 2048:3072].view(b, kvh, hd),
                                    d[:, 3072:].view(b, kvh, hd), qw.to(dev), kw.to(dev), 1e-6, pos.to(dev),
                                    inv.to(dev), slot_mapping=slots.to(dev), k_cache=kcd, v_cache=vcd, want_k_out=False)
    od = nat.paged_attn_decode(qd, kcd, vcd, bt.to(dev), cl.to(dev), hd ** -0.5, 1024)
    torch.cuda.synchronize()
    assert torch.equal(vcd.cpu(), vc_ref), "V cache slots differ"
    kdiff = (kcd.cpu().float() - kc_ref.float()).abs().max().item()
    err = ((od.cpu().float() - o_ref.float()).abs().max() / o_ref.float().abs().max()).item()
    assert kdiff <= 0.04 and err <= 2e-2, (kdiff, err)
    print(f"[smoke] decode step ok: K-cache max|diff|={kdiff:.3g}, attention rel err={err:.3g}")
