"""Load tests/golden/ref_*.npz (made by tools/make_golden.py from the reference's own source) and rebuild, from the
stored RandomState seeds, the raw draws the reference consumed -> the noise / step-frequency tensors to inject."""
import glob
import hashlib
import os

import numpy as np

from oracle.sampling_oracle import OracleConfig, SamplingOracle

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def tick_fixtures():
    return sorted(p for p in glob.glob(os.path.join(GOLDEN_DIR, 'ref_*.npz')) if not p.endswith('ref_host.npz'))


def fixture_id(path):
    return os.path.basename(path)[4:-4]


def _regen(kind, seed, shape):
    """Must mirror oracle/jaxshim/jax/random.py (numpy legacy MT19937 streams are frozen)."""
    seed = int(seed)
    shape = tuple(int(s) for s in shape if s > 0)
    if kind == 'normal':
        return np.random.RandomState(seed).standard_normal(shape).astype(np.float32)
    if kind == 'uniform01':
        u = np.random.RandomState(seed ^ 0x2545F491).random_sample(shape).astype(np.float32)
        return np.minimum(u, np.nextafter(np.float32(1), np.float32(0)))
    if kind == 'choice':
        return np.random.RandomState(seed ^ 0x1B873593).randint(0, 3, size=shape).astype(np.int32)
    raise ValueError(kind)


class Fixture:
    def __init__(self, path):
        z = np.load(path, allow_pickle=False)
        self.z = z
        self.path = path
        self.method, self.param = str(z['method']), str(z['param'])
        self.robot, self.gait = str(z['robot']), str(z['gait'])
        self.adaptive = bool(int(z['adaptive']))
        self.N = int(z['N'])
        self.mp = {k[3:]: z[k].item() for k in z.files if k.startswith('mp_')}
        draws = [(str(k), _regen(str(k), s, sh)) for k, s, sh in zip(z['draw_kinds'], z['draw_seeds'], z['draw_shapes'])]
        h = hashlib.sha256()
        for (k, raw), s in zip(draws, z['draw_seeds']):
            h.update(k.encode()); h.update(np.int64(s).tobytes()); h.update(np.ascontiguousarray(raw).tobytes())
        assert h.hexdigest() == str(z['draws_sha256']), 'regenerated draws differ from the ones the reference consumed'
        self.draws = draws

    def config_kwargs(self):
        kw = dict(sampling_method=self.method, control_parametrization=self.param, optimize_step_freq=self.adaptive,
                  num_parallel_computations=self.N)
        kw.update(self.mp)
        return kw

    def oracle(self, dtype=np.float32):
        return SamplingOracle(OracleConfig.for_robot(self.robot, dtype=dtype, **self.config_kwargs()))

    def sigma(self):
        return self.z['sigma'] if 'sigma' in self.z.files else None

    def noise(self, oracle):
        normals = [r for k, r in self.draws if k == 'normal']
        uniforms = [r for k, r in self.draws if k == 'uniform01']
        if self.method == 'random_sampling':
            assert len(normals) == 2 and np.array_equal(normals[0], normals[1])    # same key for both blocks (row a9)
            return oracle.assemble_noise_random_sampling(normals[0], uniforms[0])
        if self.method == 'mppi':
            return oracle.assemble_noise_gaussian(normals[0], np.float32(oracle.cfg.sigma_mppi))
        return oracle.assemble_noise_gaussian(normals[0], self.sigma())

    def step_freqs(self, oracle):
        if not self.adaptive:
            return None
        idx = [r for k, r in self.draws if k == 'choice'][0]
        return oracle.assemble_step_frequencies(idx, float(self.z['nominal_step_freq']), int(self.z['optimize_swing']))

    def scenario(self):
        z = self.z
        return dict(state=z['state'], reference=z['reference'], contact_sequence=z['contact_sequence'],
                    phase_signal=z['phase_signal'], nominal_step_freq=float(z['nominal_step_freq']))

    def expected(self):
        z = self.z
        out = dict(nmpc_GRFs=z['nmpc_GRFs'], nmpc_predicted_state=z['nmpc_predicted_state'],
                   best_control_parameters=z['best_control_parameters'], best_cost=z['best_cost'],
                   best_freq=z['best_freq'], costs=z['costs'], best_index=int(np.argmin(z['costs'])))
        if 'new_sigma_cem_mppi' in z.files:
            out['new_sigma_cem_mppi'] = z['new_sigma_cem_mppi']
        return out
