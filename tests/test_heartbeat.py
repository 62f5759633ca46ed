"""Failure detection: the host watchdog's logic over heartbeat counters.
The device half (a word per rank in peer-mapped HBM) is exercised in test_gpu_engine.py."""


class _Clock:
    def __init__(self):
        self.t = 0.0

    def __call__(self):
        return self.t


def test_watchdog_flags_quiet_ranks_but_not_finished_ones(dtg):
    from dtg_b200.parallel.heartbeat import Watchdog
    beats = [[0, 0], [0, 0], [0, 0]]
    clock, calls = _Clock(), []
    wd = Watchdog(lambda: [tuple(b) for b in beats], timeout_secs=5.0, on_stale=calls.append,
                  clock=clock)
    assert wd.poll() == []
    for t in range(1, 5):                      # ranks 0 and 2 progress, rank 1 is stuck
        clock.t = float(t * 2)
        beats[0][0] += 1
        beats[2][0] += 3
        stale = wd.poll()
    assert stale == [1] and calls == [[1]] and sorted(wd.alive()) == [0, 2]
    beats[2][1] = 1                            # rank 2 finishes cleanly, then stays quiet
    clock.t = 100.0
    assert wd.poll() == [0, 1]                 # 0 stopped beating too; 2 is done, not stale
    assert wd.done == [2] and calls == [[1], [0]]
    beats[1][0] += 1                           # a rank that comes back is alive again
    clock.t = 101.0
    assert wd.poll() == [0]


def test_watchdog_thread_and_hook(dtg):
    import time
    from dtg_b200.parallel.heartbeat import DeviceHeartbeatHook, Watchdog
    from dtg_b200.session.hooks import SessionRunArgs, SessionRunContext

    class FakeHeartbeat:
        def __init__(self):
            self.slots = [[0, 0], [0, 0]]

        def beat(self, n=1):
            self.slots[0][0] += n

        def mark_done(self):
            self.slots[0][1] = 1

        def read(self):
            return [tuple(s) for s in self.slots]
    hb = FakeHeartbeat()
    clock = _Clock()
    wd = Watchdog(hb.read, timeout_secs=1.0, clock=clock)
    msgs = []
    hook = DeviceHeartbeatHook(hb, wd, poll_every=2, stop_on_stale=True, log=msgs.append)
    rc = SessionRunContext(SessionRunArgs(None), None)
    for i in range(6):
        clock.t = float(i)
        hook.after_run(rc, None)
    assert hb.slots[0][0] == 6
    assert hook.reported == [1] and len(msgs) == 1 and "rank(s) [1]" in msgs[0]
    assert rc.stop_requested
    hook.end(None)
    assert hb.slots[0][1] == 1
    # background thread form
    wd2 = Watchdog(hb.read, timeout_secs=1e9).start(interval_secs=0.01)
    time.sleep(0.1)
    wd2.stop()
    assert wd2.polls >= 2 and wd2.stale == []
