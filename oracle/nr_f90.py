"""random_gen='nr_f90' of the reference (TEST INFRASTRUCTURE, like everything under oracle/): the combined
Marsaglia-shift / Park-Miller generator mars_ran (general.f90:625-676), the seeding rules of
random_seed_wrapper (general.f90:533-599) and of start.f90:383,500 / run.f90:309-311, and the two consumers
on the hot path's fixtures:

  gaunoise_vect (initcond.f90:4351-4389)       initaa='gaussian-noise' of samples/helical-MHDturb
  fconst_coefs_hel (forcing.f90:1554-1760)     the three random numbers per time step of iforce='helical'

32-bit two's-complement integer arithmetic is emulated on Python ints."""
import math

import numpy as np

_M32 = 0xFFFFFFFF
IM, IA, IQ, IR = 2147483647, 16807, 127773, 2836


def _s32(u):
    """bit pattern -> signed 32-bit value"""
    u &= _M32
    return u - (1 << 32) if u & 0x80000000 else u


class MarsRan:
    """mars_ran with its two-word state rstate (general.f90:177,625-676)"""

    def __init__(self):
        self.r1, self.r2 = 0, 0            # rstate=0 at program start
        self.init1 = 1812
        # am = nearest(1.0,-1.0)/im: the largest real below 1 over 2^31-1 (double precision build)
        self.am = math.nextafter(1.0, 0.0) / IM

    def _initialise(self, init=None):
        if init is not None:
            self.init1 = init
        a = abs(self.init1)
        self.r1 = _s32(777755555 ^ a)
        self.r2 = _s32((888889999 ^ a) | 1)

    def draw(self, init=None):
        if init is not None or self.r1 == 0 or self.r2 <= 0:
            self._initialise(init)
        u = self.r1 & _M32                                   # Marsaglia shift sequence, period 2^32-1
        u ^= (u << 13) & _M32
        u ^= u >> 17                                         # ishft(.,-17): logical shift
        u ^= (u << 5) & _M32
        self.r1 = _s32(u)
        k = self.r2 // IQ                                    # Park-Miller by Schrage's method (r2 > 0 here)
        self.r2 = IA * (self.r2 - k * IQ) - IR * k
        if self.r2 < 0:
            self.r2 += IM
        return self.am * (((self.r1 ^ self.r2) & IM) | 1)

    def draws(self, n):
        return np.array([self.draw() for _ in range(n)])

    # random_seed_wrapper(PUT=seed) for 'nr_f90' (general.f90:571-597)
    def put(self, seed):
        if seed[1] == 0:
            self.draw(init=seed[0])          # "state cannot be result from previous call, so initialize" (+ one draw)
        else:
            self.r1, self.r2 = _s32(seed[0]), _s32(seed[1])

    def get(self):
        return [self.r1, self.r2]


def start_stream(iproc=0, seed0=1812):
    """state of channel 1 when start.x begins the initial conditions (start.f90:313-385): seed=GET (0,0),
    seed(1)=-((seed0-1812+1)*10+iproc), PUT"""
    g = MarsRan()
    seed = g.get()
    seed[0] = -((seed0 - 1812 + 1) * 10 + iproc)
    g.put(seed)
    return g


def run_stream(seed0=1812):
    """state of channel 1 when run.x enters the time loop, for a var.dat written by start.x with the default
    lseed_global=T: start.f90:500-503 re-seeds with seed(1)=seed0 (seed(2) still 0 -> initialise + one draw) after
    the initial conditions, output_persistent_general stores that state in the snapshot (persist.f90:222-227), and
    run.x restores it when it reads var.dat (persist.f90:156-160; run.f90:309-311 set (seed0,seed0) before that)."""
    g = MarsRan()
    g.put([seed0, 0])
    return g


def gaunoise_vect(g, f, ampl, i1, i2):
    """initcond.f90:4351-4389 on f[var][n][m][l] incl. ghost zones: per (n,m) row and variable i, even i-i1 draws
    r(mx), p(mx) and adds ampl*sqrt(-2 ln r)*sin(2 pi p); odd i-i1 re-uses r,p with cos.  0-based i1..i2 inclusive."""
    nvar, mz, my, mx = f.shape
    for n in range(mz):
        for m in range(my):
            r = p = None
            for i in range(i1, i2 + 1):
                if (i - i1) % 2 == 0:
                    r, p = g.draws(mx), g.draws(mx)
                    tmp = np.sqrt(-2 * np.log(r)) * np.sin(2 * np.pi * p)
                else:
                    tmp = np.sqrt(-2 * np.log(r)) * np.cos(2 * np.pi * p)
                f[i, n, m, :] += ampl * tmp
    return f


def read_kdat(path):
    """k.dat: nk, kav, then kkx(nk), kky(nk), kkz(nk) in free format (forcing.f90 initialize_forcing)"""
    tok = open(path).read().split()
    nk, kav = int(tok[0]), float(tok[1])
    v = np.array(tok[2:2 + 3 * nk], dtype=np.float64)
    return nk, kav, v[:nk], v[nk:2 * nk], v[2 * nk:3 * nk]


def helical_draw(g, nk):
    """the random numbers of one call of fconst_coefs_hel (forcing.f90:1590-1594,1668): fran(2) -> phase, ik; phi"""
    fran = g.draws(2)
    phase = math.pi * (2 * fran[0] - 1.0)
    ik = int(nk * (0.9999 * fran[1])) + 1           # Fortran integer conversion truncates; 1-based
    phi = g.draw() * 2 * math.pi
    return phase, ik, phi


class MinStd:
    """random_gen='min_std' (the default, general.f90:178): ran0, the 'Minimal Standard' generator of Lewis, Goodman
    and Miller with one state word (general.f90:601-623); PUT/GET move that word (general.f90:551-555)."""
    MASK = 123459876

    def __init__(self):
        self.state = 0
        self.am = 1.0 / IM

    def draw(self):
        d = _s32(self.state ^ self.MASK)
        k = int(d / IQ)                                   # Fortran integer division truncates toward zero
        d = IA * (d - k * IQ) - IR * k
        if d < 0:
            d += IM
        self.state = _s32(d ^ self.MASK)
        return self.am * d

    def draws(self, n):
        return np.array([self.draw() for _ in range(n)])

    def put(self, seed):
        self.state = _s32(seed[0])

    def get(self):
        return [self.state, 0]


def start_stream_min_std(iproc=0, seed0=1812):
    """start.f90:383 with random_gen='min_std': rstate(1) = -((seed0-1812+1)*10+iproc), processor dependent"""
    g = MinStd()
    g.put([-((seed0 - 1812 + 1) * 10 + iproc), 0])
    return g


def run_stream_min_std(seed0=1812):
    """start.f90:500-503 puts seed(1)=seed0 after the initial conditions; the snapshot carries it into run.x"""
    g = MinStd()
    g.put([seed0, 0])
    return g
