"""Host-side logic of the symmetric slot pool (no GPU needed)."""


def test_slot_pool_tracks_ownership():
    """CPU-only logic check of SymmSlotPool: a held slot is never handed out twice."""
    from deepspeed_b200.comm.symm import SymmSlotPool

    class _Buf:
        nbytes = 64

    class _Ctx:
        def alloc(self, n):
            return _Buf()

        def wait_rs_done(self, e):
            pass

    pool = SymmSlotPool(_Ctx(), 64, n_slots=3)
    a = pool.acquire()          # the pending unit
    seen = []
    for _ in range(10):         # other units cycle
        s = pool.acquire()
        assert s is not None and s is not a
        seen.append(s)
        pool.release(s)
    assert len({id(x) for x in seen}) == 2
    b, c = pool.acquire(), pool.acquire()
    assert pool.acquire() is None and pool.n_free() == 0
    for s in (a, b, c):
        pool.release(s)
    assert pool.n_free() == 3
