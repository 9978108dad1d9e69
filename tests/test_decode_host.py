"""Host logic of the graphed greedy decoder (no GPU): the stopping rule applied after speculative generation must give
the length of the reference loop, which tests completeness after every token (captioning_bmrl_loops.py:137-150)."""
import torch

from bmhrl_b200.decode import reference_stop_length


def _reference_loop_length(tokens, max_len, end_idx):
    """The reference's loop run over a pre-drawn token matrix: tokens[:, j] is what step j would emit."""
    B = tokens.shape[0]
    n = 1
    complete = torch.zeros(B, dtype=torch.bool)
    while n <= max_len and not bool(complete.all()):
        complete |= tokens[:, n] == end_idx
        n += 1
    return n, bool(complete.all())


def test_stop_length_equals_reference_loop():
    g = torch.Generator().manual_seed(5)
    END = 3
    for case in range(300):
        B = int(torch.randint(1, 5, (1,), generator=g))
        max_len = int(torch.randint(1, 12, (1,), generator=g))
        tokens = torch.randint(2, 7, (B, max_len + 1), generator=g)
        tokens[:, 0] = 2
        n_ref, done = _reference_loop_length(tokens, max_len, END)
        # speculative generation checks only every k tokens: any look-ahead >= n_ref must give the same answer
        for n_seen in range(n_ref, max_len + 2):
            stop = reference_stop_length(tokens[:, :n_seen], END)
            assert (stop == n_ref) if done else (stop is None), (case, n_seen)
        # ... and before the last row has finished the rule must not fire
        for n_seen in range(1, n_ref):
            assert reference_stop_length(tokens[:, :n_seen], END) is None


def test_stop_length_ignores_the_start_column_and_empty_prefix():
    assert reference_stop_length(torch.tensor([[3]]), 3) is None
    assert reference_stop_length(torch.tensor([[3, 5, 3]]), 3) == 3
    assert reference_stop_length(torch.tensor([[2, 3, 3], [2, 4, 3]]), 3) == 3
    assert reference_stop_length(torch.tensor([[2, 3, 9], [2, 4, 9]]), 3) is None


def test_graphed_decoder_has_no_cpu_path(agent_sd):
    """The product path fails loudly without a CUDA device: no CPU fallback, no oracle behind it."""
    import pytest
    from conftest import PAD, build_agent
    from bmhrl_b200.decode import GraphedGreedyDecoder
    model, _ = build_agent(agent_sd, device="cpu")
    model.eval()
    model.set_inference_mode(True)
    with pytest.raises(RuntimeError, match="CUDA"):
        GraphedGreedyDecoder(model, 1, 8, 8, 5, 2, 3, PAD)
    model.train()
    with pytest.raises(RuntimeError, match="eval"):
        GraphedGreedyDecoder(model, 1, 8, 8, 5, 2, 3, PAD)
