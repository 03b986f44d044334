import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def w3d():
    import wave3d_b200
    return wave3d_b200


# Values from BASELINE.md §2.3: the reference's CUDA sources (full-precision PI) executed on a CPU
# through a <<<>>> -> loop emulation; identical for 1/2/4/8 ranks.
BASELINE_CUDA_EMULATED = {
    # (N, Lx, Ly, Lz, T, steps): {layer: (abs, rel)}
    (32, 1.0, 1.0, 1.0, 0.025, 20): {1: (2.8201634361124661e-09, 1.0969951822587042e-05),
                                     20: (1.1278897771305196e-06, 0.0043628395206666334)},
    (33, 1.0, 1.0, 1.0, 0.025, 20): {1: (2.6430055921622397e-09, 2.6520104914304403e-09),
                                     20: (1.0570376811624982e-06, 1.061135006134029e-06)},
    (40, 1.0, 2.0, 3.0, 0.05, 20): {1: (6.5632971280038532e-09, 5.7786455674841032e-05),
                                    20: (2.6241305268737136e-06, 0.027881901730471118)},
    (64, "pi", "pi", "pi", 0.025, 20): {1: (7.1502248566446269e-11, 2.4628957934303673e-06),
                                        20: (2.8600414148094444e-08, 0.0023515691066024959)},
    (128, 1.0, 1.0, 1.0, 0.025, 20): {1: (1.7624035564267615e-10, 4.6883354474921256e-05),
                                      2: (7.0496064541458736e-10, None),
                                      20: (7.0485175029944003e-08, 0.064939419919985572)},
}
