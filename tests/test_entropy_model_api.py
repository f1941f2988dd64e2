"""EntropyModel.compress / decompress argument checking and round trips, mirroring the reference's own tests
(compressai/tests/test_entropy_models.py:104-160 error cases, :258-283 compression of 0-D .. 5-D inputs)."""
import importlib

import pytest
import torch

PKG = "l-msfc-v2-training_b200"


@pytest.fixture
def em():
    return importlib.import_module(PKG + ".entropy_models")


def test_invalid_coder(em):
    with pytest.raises(ValueError):
        em.EntropyModel(entropy_coder="huffman")
    with pytest.raises(ValueError):
        em.EntropyModel(entropy_coder=0xFF)


def test_invalid_inputs(em):
    m = em.EntropyModel()
    with pytest.raises(TypeError):
        m.compress(torch.rand(1, 3))
    with pytest.raises(ValueError):
        m.compress(torch.rand(1, 3), torch.rand(2, 3))
    with pytest.raises(ValueError):
        m.compress(torch.rand(1, 3, 1, 1), torch.rand(2, 3))


def test_invalid_cdf_length_and_offsets(em):
    m = em.EntropyModel()
    x, indexes = torch.rand(1, 32, 16, 16), torch.rand(1, 32, 16, 16)
    with pytest.raises(ValueError):          # uninitialised CDFs
        m.compress(x, indexes)
    m._quantized_cdf.resize_(32, 1)
    with pytest.raises(ValueError):          # uninitialised lengths
        m.compress(x, indexes)
    m._cdf_length.resize_(32, 1)
    with pytest.raises(ValueError):          # 2-D lengths
        m.compress(x, indexes)
    m._cdf_length.resize_(32)
    with pytest.raises(ValueError):          # uninitialised offsets
        m.compress(x, indexes)


def test_invalid_decompress(em):
    m = em.EntropyModel()
    with pytest.raises(TypeError):
        m.decompress(["ssss"])
    with pytest.raises(ValueError):
        m.decompress("sss", torch.rand(1, 3, 4, 4))
    with pytest.raises(ValueError):
        m.decompress(["sss"], torch.rand(1, 4, 4))
    with pytest.raises(ValueError):
        m.decompress(["sss"], torch.rand(2, 4, 4))
    with pytest.raises(ValueError):
        m.decompress(["sss"], torch.rand(1, 4, 4), torch.rand(2, 4, 4))


def test_invalid_quantize_mode(em):
    with pytest.raises(ValueError):
        em.EntropyModel().quantize(torch.rand(1, 3), mode="toto")


@pytest.mark.gpu
def test_compression_2d(em):
    x = torch.rand(1, 128, 32, 32, device="cuda")
    eb = em.EntropyBottleneck(128).cuda()
    eb.update()
    s = eb.compress(x)
    assert isinstance(s, list) and len(s) == 1 and isinstance(s[0], bytes)
    x2 = eb.decompress(s, x.size()[2:])
    assert torch.allclose(torch.round(x), x2)


@pytest.mark.gpu
def test_compression_nd(em):
    eb = em.EntropyBottleneck(128).cuda()
    eb.update()
    x = torch.rand(1, 128, device="cuda")                      # 0-D
    assert torch.allclose(torch.round(x), eb.decompress(eb.compress(x), []))
    for i in range(1, 6):                                       # 1 .. 5 dimensions
        x = torch.rand(1, 128, *([4] * i), device="cuda")
        assert torch.allclose(torch.round(x), eb.decompress(eb.compress(x), x.size()[2:]))


@pytest.mark.gpu
def test_compress_needs_a_cuda_module(em):
    eb = em.EntropyBottleneck(8).cuda()
    eb.update()
    with pytest.raises(RuntimeError, match="(?i)cuda|no CPU fallback"):
        eb.compress(torch.rand(1, 8, 4, 4))                      # host tensor: refused, never coded on the CPU


def test_modules_pickle_and_deepcopy(em):
    """entropy_models.py:131-138 (__getstate__ / __setstate__ exist there because the pybind11 coder cannot be pickled)."""
    import copy
    import pickle
    eb = em.EntropyBottleneck(8)
    eb2 = pickle.loads(pickle.dumps(eb))
    assert eb2.entropy_coder.name == "ans" and torch.equal(eb2.quantiles, eb.quantiles)
    gc = copy.deepcopy(em.GaussianConditional([0.11, 0.5, 1.0]))
    assert gc.entropy_coder.name == "ans" and gc.scale_table.numel() == 3


@pytest.mark.gpu
def test_entropy_coder_object_speaks_the_list_api(em):
    """entropy_model.entropy_coder.encode_with_indexes(list, list, cdfs, lengths, offsets) -> the reference's bytes."""
    import os
    import numpy as np
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "rans.npz"))
    sym, idx, cdfs, sizes, offsets = (G[f"gc_outliers.{k}"] for k in ("sym", "idx", "cdfs", "sizes", "offsets"))
    m = em.EntropyModel()
    s = m.entropy_coder.encode_with_indexes(sym.tolist(), idx.tolist(), cdfs.tolist(), sizes.tolist(), offsets.tolist())
    assert s == G["gc_outliers.string"].tobytes()
    assert m.entropy_coder.decode_with_indexes(s, idx.tolist(), cdfs.tolist(), sizes.tolist(), offsets.tolist()) == sym.tolist()
