"""Row (f)3: HypergraphConv-style propagation over the CSR incidence against the float64 restatement
of torch_geometric's formula on the COO pair list (oracle_np.hypergraph_conv_propagate)."""
import numpy as np
import pytest
import torch

import hg_b200
import oracle_np as oracle
from hg_b200 import ops, synth
from test_gpu_parity import _cuda

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,d,k,f", [(600, 64, 5, 32), (5000, 128, 8, 256), (3000, 96, 16, 50), (2000, 64, 8, 3)])
def test_propagate_matches_pyg_formula(n, d, k, f):
    feats = synth.make_features(n, d, 17 + n, "gaussian")
    out = hg_b200.knn_hyperedges(_cuda(feats), k=k, kernel="gaussian", lambda_h=1.0, return_csr=True)
    x = np.random.default_rng(n).standard_normal((n, f)).astype(np.float32)
    got, m = ops.hypergraph_propagate(_cuda(x), out["csr_rowptr"], out["csr_col"], return_hyperedge_features=True)
    ref, ref_m = oracle.hypergraph_conv_propagate(x, out["edge_index"].cpu().numpy(), n)
    np.testing.assert_allclose(m.cpu().numpy()[: ref_m.shape[0]], ref_m, rtol=1e-5, atol=1e-6)
    assert (m.cpu().numpy()[ref_m.shape[0]:] == 0).all()              # ids beyond the largest hyperedge: empty
    np.testing.assert_allclose(got.cpu().numpy(), ref, rtol=1e-5, atol=1e-6)


def test_propagate_tiny_by_hand():
    # pairs (node, hyperedge): (0,1) (0,2) (1,2) (3,4); node 2 and 4 are only hyperedges, node 3 has one
    ei = torch.tensor([[0, 0, 1, 3], [1, 2, 2, 4]], dtype=torch.int64).cuda()
    rowptr, col, _ = ops.csr_from_coo(ei, 5)
    x = torch.tensor([[1.0, 2.0], [3.0, 4.0], [5.0, 6.0], [7.0, 8.0], [9.0, 10.0]]).cuda()
    out, m = ops.hypergraph_propagate(x, rowptr, col, return_hyperedge_features=True)
    # m_1 = x0, m_2 = (x0 + x1) / 2, m_4 = x3;  out_0 = (m_1 + m_2) / 2, out_1 = m_2, out_3 = m_4, others 0
    exp_m = torch.tensor([[0, 0], [1, 2], [2, 3], [0, 0], [7, 8]], dtype=torch.float32)
    exp_o = torch.tensor([[1.5, 2.5], [2, 3], [0, 0], [7, 8], [0, 0]], dtype=torch.float32)
    assert torch.equal(m.cpu(), exp_m) and torch.equal(out.cpu(), exp_o)
    with pytest.raises(ValueError):
        ops.hypergraph_propagate(x, rowptr[:-1], col)
