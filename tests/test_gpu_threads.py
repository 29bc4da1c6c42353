"""The C ABI from several host threads at once (the reference's decode() is re-entrant; the library takes one
lock per device, not per process): every thread must get exactly what a lone caller gets, whatever the interleaving,
through every entry point family (open/read/close, one-call decode, batch), on one device and — where the box has
more — on two."""
import threading

import numpy as np
import pytest

from oracle import oracle
from tools import flacgen as fg

pytestmark = pytest.mark.gpu


def _worker(z, streams, refs, device, rounds, errors, tag):
    try:
        for r in range(rounds):
            for k, (s, ref) in enumerate(zip(streams, refs)):
                which = (r + k + tag) % 3
                if which == 0:
                    got = z.decode(s, device=device).samples
                elif which == 1:
                    got = np.empty(ref.samples.size, dtype=ref.samples.dtype)
                    z.decode_into(s, got, device=device)
                else:
                    got = z.decode_batch([s, s], device=device)[1].samples
                if not np.array_equal(got, ref.samples):
                    errors.append(f"thread {tag} round {r} stream {k} path {which}: PCM differs")
    except Exception as e:  # noqa: BLE001
        errors.append(f"thread {tag}: {type(e).__name__} {e}")


@pytest.mark.parametrize("n_threads", [2, 4])
def test_concurrent_callers(n_threads):
    import torch
    import zflac_b200 as z
    streams = []
    for i in range(3):
        p = fg.c2(1.0 + 0.5 * i)
        p.seed = 900 + i
        streams.append(fg.generate(p, want_pcm=False).flac)
    streams.append(fg.generate(fg.c3(0.5), want_pcm=False).flac)
    refs = [oracle.decode(s) for s in streams]
    n_dev = torch.cuda.device_count()
    errors = []
    th = [threading.Thread(target=_worker, args=(z, streams, refs, t % n_dev, 3, errors, t)) for t in range(n_threads)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errors, errors[:5]
    assert torch.cuda.current_device() == 0  # the library put the caller's device back


def test_trim_hands_idle_blocks_back_and_decoding_goes_on():
    """zflacb200_trim: the block cache gives its idle blocks back to the driver; the next call allocates afresh."""
    import torch
    import zflac_b200 as z
    p = fg.c2(20.0)
    s = fg.generate(p, want_pcm=True)
    assert np.array_equal(z.decode(s.flac).samples, s.pcm)
    torch.cuda.synchronize()
    free_cached, _ = torch.cuda.mem_get_info()
    z.trim()
    free_trimmed, _ = torch.cuda.mem_get_info()
    assert free_trimmed > free_cached  # the PCM / scratch blocks of the call above went back
    assert np.array_equal(z.decode(s.flac).samples, s.pcm)
