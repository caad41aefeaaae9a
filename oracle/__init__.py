"""CPU oracle for the audio-ssl upstream input path.  TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, the arithmetic of the reference's log-mel front
end and two-view augmentation so that the CUDA product path can be checked against
it.  Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it; nothing under
``audio-ssl_b200/`` does, and the product path has no CPU fallback.

Parity status:
  * ``oracle.views``    -- PINNED: checked bit-for-bit against the reference's own
    ``src/augmentations`` (imported read-only from /root/reference when present) and
    against the committed fixtures in ``tests/golden/`` generated from it.
  * ``oracle.frontend`` -- the reference delegates to librosa==0.8.1 (requirements.txt:71),
    which is not vendored and not installable here.  The restatement follows librosa
    0.8.1's published algorithm and is cross-checked against torchaudio's Slaney
    mel/STFT and analytic known answers; against librosa itself parity is UNPINNED.
"""
from . import frontend, views  # noqa: F401
