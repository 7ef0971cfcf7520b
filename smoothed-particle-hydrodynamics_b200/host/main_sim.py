"""main_sim.py — the muscle-signal generator of the reference (/root/reference/src/main_sim.py:4-53)
restated for Python 3 (`n // 2`, lists instead of py2 `map` iterators; numpy optional).

muscle_simulation().run() returns 96 floats: two travelling sine waves sampled at 12 positions, each
sample doubled to 24, concatenated [w1, w2, w2, w1]; `t` advances by 1.0 per call, phase velocity 1e-4.
Parity: PINNED to the reference's own script — tests/golden/main_sim_kat.json holds the 96 doubles of calls
0, 1, 2, 500 and 1000 produced by /root/reference/src/main_sim.py itself (run under Python 3 after two mechanical
py2 -> py3 rewrites of an in-memory copy, tests/golden/make_main_sim_kat.py); tests/test_host.py compares exactly.
"""
import math


def _linspace(start, stop, num):
    # numpy.linspace semantics: start + i*step with step = (stop-start)/(num-1), last sample = stop
    if num == 1:
        return [float(start)]
    step = (stop - start) / (num - 1)
    out = [start + i * step for i in range(num)]
    out[-1] = float(stop)
    return out


def parallel_waves(n=24, time=0, phi=math.pi, amplitude=1, velocity=0.0001):
    if n % 2 != 0:
        raise NotImplementedError("Currently only supports even number of muscles!")
    j = n // 2
    row_positions = _linspace(0, 1.5 * 2 * math.pi, j)
    wave_1 = [math.sin(r - velocity * time) for r in row_positions]
    wave_2 = [math.sin(r + math.pi - velocity * time) for r in row_positions]
    normalize_sine = lambda x: (x + 1) / 2
    wave_1 = [normalize_sine(x) for x in wave_1]
    wave_2 = [normalize_sine(x) for x in wave_2]
    double_wave_1, double_wave_2 = [], []
    for i in wave_1:
        double_wave_1.append(i)
        double_wave_1.append(i)
    for i in wave_2:
        double_wave_2.append(i)
        double_wave_2.append(i)
    return (double_wave_1, double_wave_2)


class muscle_simulation():
    def __init__(self, increment=1.0):
        self.increment = increment
        self.t = 0

    def run(self, do_plot=True):
        self.contraction_array = parallel_waves(time=self.t)
        self.t += self.increment
        w1, w2 = self.contraction_array
        return [float(x) for x in (w1 + w2 + w2 + w1)]
