"""Bucket construction of the fused parameter-server push (pure host logic of parallel/fused_ps.py)."""
from singa_b200.parallel.fused_ps import make_buckets, shard_bounds

# MNIST MLP: six inner-product layers laid out back to back in the arena (sizes in 1024-element chunks)
MNIST = [("fc1", 0, 1917), ("fc2", 1917, 4886), ("fc3", 6803, 2932), ("fc4", 9735, 1467), ("fc5", 11202, 490),
         ("fc6", 11692, 6)]


def test_default_bucket_gives_two_pushes_for_the_mnist_mlp():
    b = make_buckets(MNIST, 6_000_000)
    assert b == [["fc6", "fc5", "fc4", "fc3", "fc2"], ["fc1"]]
    assert [x[-1] for x in b] == ["fc2", "fc1"]              # trigger = lowest layer of each bucket


def test_bucket_size_one_is_per_layer_and_huge_is_single():
    assert make_buckets(MNIST, 1) == [[n] for n, _, _ in reversed(MNIST)]
    assert make_buckets(MNIST, 10 ** 12) == [[n for n, _, _ in reversed(MNIST)]]


def test_every_layer_in_exactly_one_contiguous_bucket():
    for size in (1, 500_000, 2_000_000, 6_000_000, 10 ** 9):
        b = make_buckets(MNIST, size)
        flat = [n for x in b for n in x]
        assert sorted(flat) == sorted(n for n, _, _ in MNIST) and len(flat) == len(MNIST)
        rng = {n: (lo, c) for n, lo, c in MNIST}
        for x in b:                                          # contiguous chunk range, descending layer order
            for a, nxt in zip(x, x[1:]):
                assert rng[nxt][0] + rng[nxt][1] == rng[a][0]


def test_gap_in_the_arena_closes_a_bucket():
    layers = [("a", 0, 10), ("b", 10, 10), ("c", 40, 10), ("d", 50, 10)]       # hole between b and c
    assert make_buckets(layers, 10 ** 9) == [["d", "c"], ["b", "a"]]


def test_shard_bounds_cover_the_range():
    for n in (0, 1, 7, 8, 1000, 12345):
        for world in (1, 2, 4, 8):
            pieces = shard_bounds(n, world)
            assert len(pieces) == world and pieces[0][0] == 0 and pieces[-1][1] == n
            assert all(a <= b for a, b in pieces) and all(p[1] == q[0] for p, q in zip(pieces, pieces[1:]))


def test_first_layer_alone_keeps_the_exposed_push_small():
    # without the rule a huge bucket swallows fc1; with it fc1 is pushed on its own whatever the bucket size
    assert make_buckets(MNIST, 10 ** 12, first_alone=True) == [["fc6", "fc5", "fc4", "fc3", "fc2"], ["fc1"]]
    assert make_buckets(MNIST, 3_000_000, first_alone=True) == [["fc6", "fc5", "fc4", "fc3"], ["fc2"], ["fc1"]]
    assert make_buckets([("only", 0, 5)], 10, first_alone=True) == [["only"]]
