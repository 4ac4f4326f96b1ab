"""A host model of the flag-in-data neighbour exchange (hypre_b200/csrc/p2p.cuh, spmv.cuh: halo_put_phase,
halo_warp_wait / ll_read, p2p_consumer_done), run with threads as ranks under random delays.

What the device does per kernel on a channel, and what the model repeats step for step:
  put     seq = seq_put + 1; if seq > 2, wait for the acknowledgement (>= seq - 2) of every receiver marked
          out_wait_ack; store every element as (value, seq) into the receiver's buffer of parity seq & 1;
          seq_put = seq
  read    f = seq_wait + 1; spin on each needed element of in_buf[f & 1] until it carries f
  done    acknowledge f to the senders marked in_send_ack; seq_wait = f
Kernels of one rank run in stream order; every rank runs the same program (sequence of channels).

The claim under test is the flow control: buffers are double buffered and NOTHING is acknowledged between ranks that
exchange in both directions -- a sender can only reach exchange s + 2 after it received the peer's data of s + 1,
which the peer sends after its reader of s ended.  Only a receiver that does not send to its sender acknowledges.
The model fails if a reader ever finds an element overwritten by a later exchange (flag > f), reads a wrong
payload, or spins past a deadline (deadlock).  A second test switches the acknowledgements of one-directional pairs
off and expects the model to catch the overwrite -- the checker has teeth.

This tests the protocol's logic, not the memory ordering of the hardware (an element is one Python object, replaced
atomically, like the single-copy-atomic halves of an LLVal)."""
import random
import threading
import time

import pytest

DEADLINE = 20.0


class Channel:
    """one direction of an exchange for all ranks: sends[r] = [(dst, n_elements), ...]"""

    def __init__(self, nranks, sends, use_acks=True):
        self.n = nranks
        self.sends = sends
        self.recvs = [[] for _ in range(nranks)]                       # recvs[r] = [(src, n, offset), ...]
        for src in range(nranks):
            for dst, cnt in sends[src]:
                off = sum(c for _, c, _ in self.recvs[dst])
                self.recvs[dst].append((src, cnt, off))
        self.buf = [[[(0.0, 0)] * sum(c for _, c, _ in self.recvs[r]) for _ in range(2)] for r in range(nranks)]
        self.seq_put = [0] * nranks
        self.seq_wait = [0] * nranks
        self.ack = {}                                                   # (sender, receiver) -> last sequence consumed
        self.wait_ack = {}
        for src in range(nranks):
            for dst, _ in sends[src]:
                one_way = all(d != src for d, _ in sends[dst])          # dst does not send to src
                self.wait_ack[(src, dst)] = one_way and use_acks
                self.ack[(src, dst)] = 0


def payload(chan_id, src, seq, k):
    return float(((chan_id * 131 + src) * 1009 + seq) * 4099 + k)


def kernel(rank, cid, ch, rng, errors, slow):
    t0 = time.monotonic()
    # ---- put phase
    seq = ch.seq_put[rank] + 1
    if seq > 2:
        for dst, _ in ch.sends[rank]:
            if ch.wait_ack[(rank, dst)]:
                while ch.ack[(rank, dst)] < seq - 2:
                    if time.monotonic() - t0 > DEADLINE:
                        errors.append("rank %d: no acknowledgement from %d (seq %d)" % (rank, dst, seq))
                        return False
                    time.sleep(0)
    for dst, cnt in ch.sends[rank]:
        off = next(o for s, _, o in ch.recvs[dst] if s == rank)
        b = ch.buf[dst][seq & 1]
        for k in range(cnt):
            b[off + k] = (payload(cid, rank, seq, k), seq)
    ch.seq_put[rank] = seq
    # ---- rows: interior work, then the elements this rank needs (every received element is needed by some row)
    if rank in slow:
        time.sleep(rng.random() * 0.004)
    f = ch.seq_wait[rank] + 1
    b = ch.buf[rank][f & 1]
    for src, cnt, off in ch.recvs[rank]:
        order = list(range(cnt))
        rng.shuffle(order)
        for k in order:
            while True:
                v, flag = b[off + k]
                if flag == f:
                    break
                if flag > f:
                    errors.append("rank %d: element %d of %d overwritten by exchange %d before exchange %d was read"
                                  % (rank, k, src, flag, f))
                    return False
                if time.monotonic() - t0 > DEADLINE:
                    errors.append("rank %d: element %d of %d never arrived (seq %d)" % (rank, k, src, f))
                    return False
                time.sleep(0)
            if v != payload(cid, src, f, k):
                errors.append("rank %d: wrong payload from %d" % (rank, src))
                return False
        if rank in slow:
            time.sleep(rng.random() * 0.002)
    # ---- consumer done
    for src, _, _ in ch.recvs[rank]:
        if ch.wait_ack[(src, rank)]:
            ch.ack[(src, rank)] = f
    ch.seq_wait[rank] = f
    return True


def run(nranks, channels, program, seed, slow=()):
    errors = []

    def body(rank):
        rng = random.Random(seed * 1000 + rank)
        for cid in program:
            if errors or not kernel(rank, cid, channels[cid], rng, errors, slow):
                return
            if rng.random() < 0.2:
                time.sleep(rng.random() * 0.001)

    threads = [threading.Thread(target=body, args=(r,)) for r in range(nranks)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(DEADLINE * 3)
    assert not any(t.is_alive() for t in threads), "model did not finish"
    return errors


def symmetric_channel(nranks, rng, use_acks=True):
    """every pair exchanges in both directions (the operators A_l: structurally symmetric)"""
    sends = [[] for _ in range(nranks)]
    for a in range(nranks):
        for b in range(a + 1, nranks):
            if rng.random() < 0.6 or b == a + 1:
                sends[a].append((b, rng.randint(1, 6)))
                sends[b].append((a, rng.randint(1, 6)))
    return Channel(nranks, sends, use_acks)


def one_way_channel(nranks, rng, use_acks=True):
    """some pairs exchange in one direction only (possible for the rectangular P_l)"""
    sends = [[] for _ in range(nranks)]
    for a in range(nranks):
        for b in range(nranks):
            if a != b and rng.random() < 0.4:
                sends[a].append((b, rng.randint(1, 5)))
    if not any(sends):
        sends[0].append((nranks - 1, 3))
    return Channel(nranks, sends, use_acks)


@pytest.mark.parametrize("nranks,seed", [(2, 1), (3, 2), (4, 3), (8, 4)])
def test_exchange_needs_no_acknowledgement_between_two_way_neighbours(nranks, seed):
    rng = random.Random(seed)
    channels = [symmetric_channel(nranks, rng) for _ in range(3)]
    assert not any(any(ch.wait_ack.values()) for ch in channels)        # nothing is acknowledged at all
    program = [rng.randrange(3) for _ in range(120)]
    errors = run(nranks, channels, program, seed, slow={seed % nranks})
    assert not errors, errors[:3]


@pytest.mark.parametrize("nranks,seed", [(2, 5), (4, 6), (8, 7)])
def test_one_way_neighbours_are_acknowledged_per_pair(nranks, seed):
    rng = random.Random(seed)
    channels = [symmetric_channel(nranks, rng), one_way_channel(nranks, rng), one_way_channel(nranks, rng)]
    assert any(any(ch.wait_ack.values()) for ch in channels[1:])
    program = [rng.randrange(3) for _ in range(120)]
    errors = run(nranks, channels, program, seed, slow={0, nranks - 1})
    assert not errors, errors[:3]


def test_model_catches_a_missing_acknowledgement():
    # rank 0 sends to rank 1, which never sends back and is slow: without the acknowledgement rank 0 runs ahead
    # and overwrites the buffer of exchange s with s + 2 before rank 1 has read it
    ch = Channel(2, [[(1, 4)], []], use_acks=False)
    errors = run(2, [ch], [0] * 60, seed=11, slow={1})
    assert errors and "overwritten" in errors[0], errors[:2]
    ch = Channel(2, [[(1, 4)], []], use_acks=True)
    assert not run(2, [ch], [0] * 60, seed=11, slow={1})
