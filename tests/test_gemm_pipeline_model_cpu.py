"""Executable model of the synchronisation protocol of the tensor-core GEMM kernels (csrc/tc_gemm.cu "TS" kernel and
csrc/tc_gemm_streamk.inl): four roles (TMA producer, converters, MMA issuer, drain) talking through mbarriers with
phase-parity waits, a ring of stages and two accumulator buffers.  The loops below restate the kernels' counter
arithmetic (stage / iteration, global chunk counter, first/last-in-chunk) line by line; a randomised scheduler runs them
with asynchronous TMA completions and tcgen05.commit arrivals, and shadow state checks what the hardware would not
forgive: a stage overwritten before the MMAs that read it retired, an MMA on a stage that is not converted yet, an
accumulator buffer reused before it was drained, a parity wait that lags two phases behind, a deadlock.

This is a model, not the device code — it cannot catch a typo in the PTX — but it pins the protocol the stream-K
kernel relies on when one CTA walks several (tile, K-range) segments through the same barriers."""
import random

import pytest

STAGES, CHUNK = 4, 4


class MBar:
    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0

    def arrive(self):
        self.pending -= 1
        assert self.pending >= 0, "more arrivals than the barrier expects in one phase"
        if self.pending == 0:
            self.phase += 1
            self.pending = self.count

    def passed(self, parity, waiter_phase):
        # try_wait.parity: true once the phase with this parity has completed.  The waiter must not be two phases
        # behind (the parity would alias) — `waiter_phase` is how many completions the waiter has consumed so far.
        assert self.phase - waiter_phase <= 1, "a parity wait lags two phases behind its barrier"
        return (self.phase & 1) != parity


class Model:
    def __init__(self, segments, rng, n_conv=2, n_drain=2):
        self.segments = segments            # list of K-block counts, one per segment of this CTA
        self.rng = rng
        self.full = [MBar(1) for _ in range(STAGES)]
        self.conv = [MBar(n_conv) for _ in range(STAGES)]
        self.empty = [MBar(1) for _ in range(STAGES)]
        self.acc_full = [MBar(1) for _ in range(2)]
        self.acc_empty = [MBar(n_drain) for _ in range(2)]
        self.stage = ["free"] * STAGES      # free -> loading -> raw -> converted (per converter) -> free
        self.stage_kb = [None] * STAGES
        self.conv_done = [0] * STAGES
        self.acc = [{"owner": "mma", "kbs": [], "readers": 0} for _ in range(2)]
        self.async_events = []              # callables that complete later (TMA bytes landing, commits retiring)
        self.n_conv, self.n_drain = n_conv, n_drain
        self.drained = []                   # per drain thread: list of (segment, [K blocks]) it accumulated
        self.mma_inflight = []              # (stage) reads not retired yet

    # ---- roles: generators that yield a zero-argument predicate to wait on --------------------------------------
    def producer(self):
        st = it = 0
        kb_global = 0
        for seg, nkb in enumerate(self.segments):
            for i in range(nkb):
                yield lambda st=st, it=it: self.empty[st].passed((it & 1) ^ 1, it)
                assert self.stage[st] == "free", f"TMA overwrites stage {st} in state {self.stage[st]}"
                self.stage[st], self.stage_kb[st] = "loading", (seg, i)

                def land(st=st):
                    self.stage[st] = "raw"
                    self.full[st].arrive()
                self.async_events.append(land)
                kb_global += 1
                st += 1
                if st == STAGES:
                    st, it = 0, it + 1

    def converter(self, who):
        st = it = 0
        for _ in range(sum(self.segments)):
            yield lambda st=st, it=it: self.full[st].passed(it & 1, it)
            assert self.stage[st] == "raw" or (self.stage[st] == "converting"), self.stage[st]
            self.stage[st] = "converting"
            yield lambda: True                                   # the conversion takes a while
            self.conv_done[st] += 1
            if self.conv_done[st] == self.n_conv:
                self.stage[st], self.conv_done[st] = "converted", 0
            self.conv[st].arrive()
            st += 1
            if st == STAGES:
                st, it = 0, it + 1

    def mma(self):
        st = it = 0
        gc = 0
        for seg, nkb in enumerate(self.segments):
            for i in range(nkb):
                buf, use = gc & 1, gc >> 1
                first = i % CHUNK == 0
                last = i % CHUNK == CHUNK - 1 or i == nkb - 1
                if first:
                    yield lambda buf=buf, use=use: self.acc_empty[buf].passed((use & 1) ^ 1, use)
                    a = self.acc[buf]
                    assert a["owner"] == "mma" and a["readers"] == 0, f"accumulator {buf} reused before it was drained"
                    a["kbs"] = []
                yield lambda st=st, it=it: self.conv[st].passed(it & 1, it)
                assert self.stage[st] == "converted", f"MMA on stage {st} in state {self.stage[st]}"
                assert self.stage_kb[st] == (seg, i), "MMA consumed a stage holding another K block"
                self.acc[buf]["kbs"].append((seg, i))
                self.stage[st] = "reading"

                def retire(st=st, buf=buf, last=last):
                    self.stage[st] = "free"
                    self.empty[st].arrive()
                    if last:
                        self.acc[buf]["owner"] = "drain"
                        self.acc[buf]["readers"] = self.n_drain
                        self.acc_full[buf].arrive()
                # commits retire in issue order: chain them
                self.mma_inflight.append(retire)
                if len(self.mma_inflight) == 1:
                    self.async_events.append(self._retire_next)
                if last:
                    gc += 1
                st += 1
                if st == STAGES:
                    st, it = 0, it + 1

    def _retire_next(self):
        self.mma_inflight.pop(0)()
        if self.mma_inflight:
            self.async_events.append(self._retire_next)

    def drain(self, who):
        gc = 0
        mine = []
        self.drained.append(mine)
        for seg, nkb in enumerate(self.segments):
            got = []
            for _c in range((nkb + CHUNK - 1) // CHUNK):
                buf, use = gc & 1, gc >> 1
                yield lambda buf=buf, use=use: self.acc_full[buf].passed(use & 1, use)
                a = self.acc[buf]
                assert a["owner"] == "drain", "drain read an accumulator the MMAs are still writing"
                got += a["kbs"]
                a["readers"] -= 1
                if a["readers"] == 0:
                    a["owner"] = "mma"
                self.acc_empty[buf].arrive()
                gc += 1
            mine.append((seg, got))
            yield lambda: True                                   # the epilogue (stores / red.adds) takes a while

    # ---- scheduler ----------------------------------------------------------------------------------------------
    def run(self):
        roles = [self.producer(), self.mma()] + [self.converter(i) for i in range(self.n_conv)] + \
                [self.drain(i) for i in range(self.n_drain)]
        waiting = {}
        for r in roles:
            try:
                waiting[r] = next(r)
            except StopIteration:
                pass
        steps = 0
        while waiting or self.async_events:
            steps += 1
            assert steps < 2_000_000, "model did not terminate"
            runnable = [r for r, cond in waiting.items() if cond()]
            choices = len(runnable) + len(self.async_events)
            assert choices > 0, "deadlock: every role waits and nothing is in flight"
            k = self.rng.randrange(choices)
            if k < len(runnable):
                r = runnable[k]
                try:
                    waiting[r] = r.send(None)
                except StopIteration:
                    del waiting[r]
            else:
                # TMA completions may land out of order; commits are chained and therefore in order
                self.async_events.pop(k - len(runnable))()
        return self


def _check(m):
    for mine in m.drained:
        assert [seg for seg, _ in mine] == list(range(len(m.segments)))
        for seg, got in mine:
            assert got == [(seg, i) for i in range(m.segments[seg])], "a segment's accumulator missed or mixed K blocks"
    assert all(s == "free" for s in m.stage)


@pytest.mark.parametrize("segments", [[1], [3], [4], [5], [50], [13, 50, 24], [36, 50], [1, 1, 1, 1, 1, 1, 1], [2, 9, 4, 4, 17],
                                      [3] * 21, [8, 8, 8], [7, 1, 50]])
def test_pipeline_protocol_holds_under_random_schedules(segments):
    for seed in range(12):
        _check(Model(segments, random.Random(seed)).run())


def test_single_segment_is_the_one_tile_kernel():
    # tc_gemm_ts_kernel is the special case of one segment per CTA (8 + 8 converter / drain warps = arrival counts)
    _check(Model([50], random.Random(0), n_conv=8, n_drain=8).run())


def test_the_model_notices_a_broken_protocol():
    class Broken(Model):
        def producer(self):                                    # waits with the wrong parity: overwrites live stages
            st = it = 0
            for seg, nkb in enumerate(self.segments):
                for i in range(nkb):
                    yield lambda: True
                    assert self.stage[st] == "free", "TMA overwrites a live stage"
                    self.stage[st], self.stage_kb[st] = "loading", (seg, i)

                    def land(st=st):
                        self.stage[st] = "raw"
                        self.full[st].arrive()
                    self.async_events.append(land)
                    st += 1
                    if st == STAGES:
                        st, it = 0, it + 1
    with pytest.raises(AssertionError):
        for seed in range(20):
            Broken([20], random.Random(seed)).run()
