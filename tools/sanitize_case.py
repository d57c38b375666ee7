"""The case the compute-sanitizer runs cover (memcheck / racecheck / synccheck, one tool per run):
a small population through build, training steps (tcgen05 forward / dgrad / wgrad, BatchNorm with the fused finalize,
pooling, losses, Adam on twelve streams), validation, test evaluation, replication + readout, and the autoencoder
decode (tcgen05 GEMM + recheck).    compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / 'tests'))
import numpy as np
import parity_util as U
from serann_b200.codec import RibosomalAutoencoder
from serann_b200.synthetic import synthetic_decoder_weights

srcs, irs = U.small_population(n_generated=2)
data = U.small_dataset(n_train=400, n_test=64)
eng, _ = U.make_engine_and_oracle(irs, data, max_batch=160)
res = eng.fit(1, 160, 80, shuffle_seed=1)                 # 2 steps at 160 rows (multi-chunk BN, split weight gradients) + validation
eng.evaluate(1, 0, 64, 32)
probs, packed, mut = eng.replicate(data['g_test'][:len(irs)], 3)
assert np.isfinite(res['train_loss']).any()
eng.close()
ae = RibosomalAutoencoder(350, synthetic_decoder_weights())
tok = ae.decode_to_sequence(np.random.default_rng(0).integers(0, 2, (300, 100)))
assert tok.shape == (300, 350)
ae.close()
print('sanitize case ok')
