"""Ledger rules (DistSys/block.go, blockData.go, blockchain.go, honest.go:631-653)."""
import numpy as np
import pytest

from biscotti_b200.ledger.chain import Blockchain, UpdateRecord, go_float


def test_genesis_and_empty_blocks_are_identical_across_peers():
    a, b = Blockchain(5), Blockchain(5)
    assert len(a) == 1 and a.latest_iteration == -1
    assert a.latest_hash() == b.latest_hash()
    assert np.all(a.weights() == 0)
    # empty blocks carry Timestamp 0 => every peer mints the same one (block.go:32-38)
    ha = a.add_block(0, np.zeros(5), [], {0: 10})
    hb = b.add_block(0, np.zeros(5), [], {0: 10})
    assert ha == hb and a == b


def test_hash_links_and_verify_and_tamper():
    c = Blockchain(3)
    u = UpdateRecord(source=2, iteration=0, commitment=bytes(range(64)), signatures=[b"s" * 64])
    c.add_block(0, [0.1, -0.2, 0.3], [u], {0: 10, 2: 15}, timestamp=7)
    c.add_block(1, [0.2, -0.1, 0.3], [], None)
    assert c.verify()
    assert c.prev_hash(2) == c.block_hash(1)
    info = c.block_info(1)
    assert info["iteration"] == 0 and info["timestamp"] == 7 and info["stake"] == {0: 10, 2: 15}
    assert info["updates"][0]["source"] == 2 and info["updates"][0]["nsigs"] == 1
    assert c.get_block_index(0) == 1 and c.get_block_index(5) == -1


def test_serialize_round_trip_and_longest_chain():
    c = Blockchain(4)
    for it in range(3):
        c.add_block(it, np.arange(4) * 0.5 + it, [UpdateRecord(1, it, b"\x01" * 64, delta=np.ones(4))], {1: 10 + 5 * it},
                    timestamp=it + 1)
    d = Blockchain.deserialize(c.serialize())
    assert d == c and d.verify() and d.dump() == c.dump()
    with pytest.raises(ValueError):
        Blockchain.deserialize(b"garbage")


def test_block_gossip_accept_and_better_block_replacement():
    a, b = Blockchain(2), Blockchain(2)
    a.add_block(0, [1.0, 2.0], [UpdateRecord(3, 0, b"\x02" * 64)], {3: 15}, timestamp=9)
    assert b.accept_block(a.serialize_block(-1)) == 1          # appended
    assert b.accept_block(a.serialize_block(-1)) == 0          # duplicate
    assert a == b
    # c minted an empty block for the same height; a non-empty one replaces it (honest.go:631-653)
    c = Blockchain(2)
    c.add_block(0, [0.0, 0.0], [], None)
    assert c.accept_block(a.serialize_block(-1)) == 2
    assert c == a
    blk = bytearray(a.serialize_block(-1))
    blk[-3] ^= 1
    assert Blockchain(2).accept_block(bytes(blk)) < 0          # tampered payload: hash mismatch


def test_print_chain_format_and_go_floats():
    c = Blockchain(2)
    c.add_block(0, [0.5, 1e6], [UpdateRecord(1, 0, b"\x00" * 64)], None, timestamp=1)
    text = c.dump()
    assert text.startswith("Prev. hash: \nData: Iteration: -1, GlobalW: [0,0], deltas: []\nHash: ")
    assert "Iteration: 0, GlobalW: [0.5,1e+06]" in text
    assert "Commitment:bn256.G1:(" in text
    assert go_float(0.1) == "0.1" and go_float(123456789.0) == "1.23456789e+08"
    assert go_float(1e-5) == "1e-05" and go_float(-2.5) == "-2.5" and go_float(100000.0) == "100000"


def test_add_commit_block_is_byte_identical_to_update_records():
    """The array-based fast path of the engine's common block must hash exactly like the UpdateRecord path."""
    import numpy as np

    from biscotti_b200.ledger.chain import Blockchain, UpdateRecord

    rs = np.random.RandomState(3)
    w = rs.randn(12)
    srcs = [2, 5, 9]
    commits = [bytes(rs.randint(0, 256, 64).astype(np.uint8)) for _ in srcs]
    sigs = [[bytes(rs.randint(0, 256, 64).astype(np.uint8)) for _ in range(2)] for _ in srcs]
    stake = [10 + i for i in range(6)]
    a, b = Blockchain(12), Blockchain(12)
    ha = a.add_block(0, w, [UpdateRecord(source=s, iteration=0, commitment=c, signatures=g) for s, c, g in zip(srcs, commits, sigs)],
                     stake, timestamp=1)
    hb = b.add_commit_block(0, w, srcs, b"".join(commits), b"".join(x for g in sigs for x in g), 2, stake, timestamp=1)
    assert ha == hb and a == b and a.dump() == b.dump()
    # without signatures
    ha = a.add_block(1, w, [UpdateRecord(source=s, iteration=1, commitment=c) for s, c in zip(srcs, commits)], stake, timestamp=2)
    hb = b.add_commit_block(1, w, srcs, b"".join(commits), b"", 0, stake, timestamp=2)
    assert ha == hb and a.serialize() == b.serialize()


def test_append_packed_is_byte_identical_and_credits_stake():
    """The engine's one-call block tail (rows of commitment || approvals picked by peer id, stake +5 per contributor)
    must produce exactly the block the UpdateRecord path produces."""
    import numpy as np

    from biscotti_b200.ledger.chain import Blockchain, UpdateRecord

    rs = np.random.RandomState(4)
    w = rs.randn(12)
    n, k = 10, 3
    pack = rs.randint(0, 256, (n, 64 * (1 + k))).astype(np.uint8)
    srcs = [1, 4, 7, 8]
    stake_a = [10 + i for i in range(n)]
    stake_b = list(stake_a)
    a, b = Blockchain(12), Blockchain(12)
    for s in srcs:
        stake_a[s] += 5
    ups = [UpdateRecord(source=s, iteration=0, commitment=pack[s, :64].tobytes(),
                        signatures=[pack[s, 64 + 64 * j: 128 + 64 * j].tobytes() for j in range(k)]) for s in srcs]
    ha = a.add_block(0, w, ups, stake_a, timestamp=1)
    hb = b.append_packed(0, w, srcs, pack, k, stake_b, 5, timestamp=1)
    assert ha == hb and a == b and stake_a == stake_b and a.dump() == b.dump()
    ha = a.add_block(1, w, [], stake_a, timestamp=0)
    hb = b.append_packed(1, w, [], pack, k, stake_b, 5, timestamp=0)          # the empty block
    assert ha == hb and a.serialize() == b.serialize()
