"""CPU oracle for the speculative-decoding verification hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``duodecoding_b200/`` may import this
package; it is used by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` as the checker and
the timed CPU baseline, never as the product path.

Parity status: PINNED.  ``oracle/gen_golden.py`` runs the unmodified reference
(imported from /root/reference in the build container) on seeded inputs and
stores inputs+outputs under ``tests/golden/``; ``tests/test_oracle_golden.py``
requires this restatement to reproduce them bit-for-bit, and
``tests/test_oracle_kat.py`` carries the known-answer vectors of the
reference's own unit tests.
"""

from .reference_port import *  # noqa: F401,F403
