"""Small invocation of every kernel family, for compute-sanitizer (one tool per gpurun call)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from types import SimpleNamespace
from radiosim_b200.jets.jet import render_batch
from radiosim_b200.jets.packer import pack_jets
from radiosim_b200.survey import create_survey
from radiosim_b200 import mojave, ppdisks
for size, n in [(256, 160), (128, 40), (100, 6), (300, 3)]:
    np.random.seed(size)
    b = pack_jets(n, size, (3, 10), "normalize")
    img, m, p, _ = render_batch(b, True, planes=(size == 100), f64=False)
    assert np.isfinite(img).all()
conf = SimpleNamespace(jet=SimpleNamespace(training_type="list", scaling="normalize", num_jet_components=[3, 10]),
                       survey=SimpleNamespace(num_sources=30, class_distribution=[2, 1, 2], scale_sources=True))
np.random.seed(1)
x, y = create_survey(np.broadcast_to(np.zeros(1), (2, 3, 200, 200)), conf)
assert np.isfinite(x).all()
for fn in (mojave.gen_compact, mojave.gen_one_jet, mojave.gen_two_jet):
    assert np.isfinite(fn(3, size=96)).all()
p, _ = ppdisks.generate_proto_set(128, 2, seed=3)
assert np.isfinite(np.array(p)).all()
print("sanitize_small: ok")
