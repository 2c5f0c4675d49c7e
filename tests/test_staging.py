"""StepStaging (host plumbing for the e2e path): layout, alignment, one-copy upload."""
import torch

from vllm_triton_backend_b200.staging import StepStaging


def test_views_are_aligned_disjoint_and_roundtrip():
    spec = [("q", (5, 4, 128), torch.bfloat16), ("slots", (5,), torch.int64), ("bt", (3, 7), torch.int32),
            ("sl", (3,), torch.int32), ("empty", (0, 8), torch.bfloat16)]
    s = StepStaging(spec, "cpu")
    assert s.payload_bytes == 5 * 4 * 128 * 2 + 5 * 8 + 3 * 7 * 4 + 3 * 4
    spans = []
    for name, shape, dtype in spec:
        h, d = s.host[name], s.dev[name]
        assert tuple(h.shape) == shape and h.dtype == dtype and tuple(d.shape) == shape and d.dtype == dtype
        assert h.data_ptr() % 16 == 0 and d.data_ptr() % 16 == 0
        spans.append((h.data_ptr(), h.data_ptr() + h.numel() * h.element_size()))
    spans.sort()
    assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))
    g = torch.Generator().manual_seed(0)
    s.host["q"].copy_(torch.rand(5, 4, 128, generator=g).bfloat16())
    s.host["slots"].copy_(torch.tensor([3, -1, 9, 2**40, 0]))
    s.host["bt"].copy_(torch.arange(21, dtype=torch.int32).view(3, 7))
    s.host["sl"].copy_(torch.tensor([1, 2, 3], dtype=torch.int32))
    s.upload()
    for name, _, _ in spec:
        assert torch.equal(s.host[name], s.dev[name])
