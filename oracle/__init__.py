"""CPU oracle for the flow-warping / deformable-sampling hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``videoframeinterpolation_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs use it, and there only as the checker / the timed CPU baseline.

Parity status: the reference ships no tests or golden vectors, so the pins
are outputs of the reference's own code run in the build container on seeded inputs
(``tests/golden/make_golden.py`` -> ``tests/golden/*.npz``); for splatting, which the reference
implements only as a CUDA kernel string, the pin is that kernel text compiled as C++
(``oracle/build_oracle.py`` -> ``oracle/_ref/``).  ``tests/test_oracle_golden.py`` checks every
oracle function against those fixtures.
"""
from .torch_oracle import *  # noqa: F401,F403
