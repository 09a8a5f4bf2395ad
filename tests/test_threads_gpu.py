"""SURVEY.md 8(b), threading: the reference's members are re-entrant on distinct objects (pure functions of their
arguments), so the C ABI must be callable from several host threads at once on DISTINCT handles.  Four threads, one
context each on the same GPU, run different calls of the path concurrently (ctypes drops the GIL inside the
library); every result is compared with the oracle, and the per-thread error text stays per thread."""
import threading

import numpy as np
import pytest

from conftest import SEED

pytestmark = pytest.mark.gpu


def test_distinct_handles_from_concurrent_host_threads(port):
    from lifeapi_b200 import E_ARG, Engine, LifeB200Error
    n_threads, rounds = 4, 6
    soups = [port.soup(1000 * t, 3000 + 37 * t, SEED, t & 1) for t in range(n_threads)]
    gens = [5, 64, 33, 200]
    want_step = [port.step_batch(soups[t], gens[t]) for t in range(n_threads)]
    want_survey = [port.survey(5000 * t, 2049, SEED, 0, 48, 2049) for t in range(n_threads)]
    errors, barrier = [], threading.Barrier(n_threads)

    def work(t):
        try:
            eng = Engine(0)
            try:
                barrier.wait(timeout=60)
                for r in range(rounds):
                    got = eng.step_batch_host(soups[t], gens[t])
                    if not (got == want_step[t]).all():
                        raise AssertionError(f"thread {t} round {r}: step_batch_host differs from the oracle")
                    hits, ps, hx = eng.soup_survey(5000 * t, 2049, SEED, 0, 48, 2049)
                    ids, pops, wps, whx = want_survey[t]
                    if not ((hits["id"] == ids).all() and (ps, hx) == (wps, whx)):
                        raise AssertionError(f"thread {t} round {r}: soup survey differs from the oracle")
                    # a failing call on this thread: its message must be this thread's own (thread-local last error)
                    try:
                        eng.set_lanes_per_pattern(3 + 2 * t)  # 3, 5, 7, 9: never valid
                        raise AssertionError("bad lane grouping accepted")
                    except LifeB200Error as e:
                        if e.code != E_ARG or "lanes per pattern" not in str(e):
                            raise AssertionError(f"thread {t}: wrong error {e}")
            finally:
                eng.close()
        except Exception as e:  # noqa: BLE001 - reported by the main thread
            errors.append((t, repr(e)))
            try:
                barrier.abort()
            except Exception:
                pass

    threads = [threading.Thread(target=work, args=(t,)) for t in range(n_threads)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=300)
    assert not any(th.is_alive() for th in threads), "a worker thread hung"
    assert not errors, errors
