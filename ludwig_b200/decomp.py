"""Cartesian domain decomposition, as cs_init does it (src/coords.c:173-215):
ranks in a periodic 3-d grid, z fastest (MPI_Cart_create order), uniform blocks.
Pure host logic: no device needed."""


def dims_create(nranks, ndim=3):
    """MPI_Dims_create-like balanced factorisation, largest first, padded with 1s."""
    dims = [1] * ndim
    n = nranks
    f = 2
    factors = []
    while n > 1:
        while n % f == 0:
            factors.append(f)
            n //= f
        f += 1
    for fac in sorted(factors, reverse=True):
        i = dims.index(min(dims))
        dims[i] *= fac
    dims.sort(reverse=True)
    return tuple(dims + [1] * (3 - ndim))


class Decomposition:
    def __init__(self, ntotal, rank, nranks, grid=None, periodic=(1, 1, 1)):
        ndim = 2 if ntotal[2] == 1 else 3
        self.cartsz = tuple(grid) if grid else dims_create(nranks, ndim)
        assert self.cartsz[0] * self.cartsz[1] * self.cartsz[2] == nranks
        self.ntotal = tuple(ntotal)
        self.rank, self.nranks = rank, nranks
        self.periodic = tuple(periodic)
        self.coords = self.rank_to_coords(rank)
        for a in range(3):
            if ntotal[a] % self.cartsz[a] != 0:
                raise ValueError("size %d not divisible by grid %d in direction %d"
                                 % (ntotal[a], self.cartsz[a], a))
        self.nlocal = tuple(ntotal[a] // self.cartsz[a] for a in range(3))
        self.offset = tuple(self.coords[a] * self.nlocal[a] for a in range(3))
        self.nbr_lo = tuple(self._nbr(a, -1) for a in range(3))
        self.nbr_hi = tuple(self._nbr(a, +1) for a in range(3))

    def rank_to_coords(self, rank):
        cz = rank % self.cartsz[2]
        cy = (rank // self.cartsz[2]) % self.cartsz[1]
        cx = rank // (self.cartsz[2] * self.cartsz[1])
        return (cx, cy, cz)

    def coords_to_rank(self, c):
        return (c[0] * self.cartsz[1] + c[1]) * self.cartsz[2] + c[2]

    def _nbr(self, a, d):
        c = list(self.coords)
        c[a] += d
        if c[a] < 0 or c[a] >= self.cartsz[a]:
            if not self.periodic[a]:
                return -1
            c[a] %= self.cartsz[a]
        return self.coords_to_rank(c)
