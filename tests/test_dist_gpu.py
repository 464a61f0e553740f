"""Multi-GPU path on real devices: world_size-2 / -3 jobs through the
npge_anchor_dist_* entry points of the C ABI, compared with the oracle on the
whole job (bit-exact).  With fewer GPUs than ranks the ranks share cuda:0 and
the collectives go through gloo; with enough GPUs they use NCCL."""
import pytest

from test_dist_cpu import launch

pytestmark = pytest.mark.gpu


EXCHANGES = ["peer", "bulk", "collective"]


@pytest.mark.parametrize("exchange", EXCHANGES)
@pytest.mark.parametrize("k", [12, 20, 27])
def test_two_ranks_tiny(k, exchange):
    launch(2, ["cuda", "tiny", 1.0, k, "interleaved"],
           extra_env={"NPGE_ANCHOR_EXCHANGE": exchange})


@pytest.mark.parametrize("exchange", EXCHANGES)
def test_three_ranks_tiny_blocks(exchange):
    launch(3, ["cuda", "tiny", 1.0, 16, "blocks"],
           extra_env={"NPGE_ANCHOR_EXCHANGE": exchange})


@pytest.mark.parametrize("exchange", EXCHANGES)
def test_two_ranks_small(exchange):
    out = launch(2, ["cuda", "small", 1.0, 20, "interleaved"],
                 extra_env={"NPGE_ANCHOR_EXCHANGE": exchange})
    assert "blocks" in out[0]


@pytest.mark.parametrize("exchange", ["bulk", "collective"])
def test_one_sequence_over_three_ranks(exchange):
    """Positions are sharded, not sequences: one chromosome over three ranks."""
    launch(3, ["cuda", "oneseq", 1.0, 20, "interleaved"],
           extra_env={"NPGE_ANCHOR_EXCHANGE": exchange})
