"""Random flat JSON objects over the sensor member names (shared by the CPU and GPU JSON tests)."""
import numpy as np

import b200_sensor as bs

KEYS = list(bs.schemas.FEATURE_NAMES)
LABEL = "failure_occurred"


def rnd_objects(seed, n):
    g = np.random.default_rng(seed)

    def num():
        t = g.integers(0, 8)
        if t == 0:
            return str(int(g.integers(-5000, 5000)))
        if t == 1:
            return repr(float(np.float32(g.uniform(-1000, 1000))))
        if t == 2:
            return "%.17g" % g.normal()
        if t == 3:
            return "%de%d" % (g.integers(-99, 99), g.integers(-30, 30))
        if t == 4:
            return "%.6fE%+d" % (g.uniform(-10, 10), g.integers(-40, 40))
        if t == 5:
            return "0"
        if t == 6:
            return "-0.0"
        return str(g.integers(0, 2 ** 63))

    def val(depth=0):
        t = g.integers(0, 10)
        if t < 5:
            return num()
        if t == 5:
            return '"s\\"x\\u00e9\\n"'
        if t == 6:
            return ["true", "false", "null"][g.integers(0, 3)]
        if t == 7 and depth < 3:
            return "[" + ",".join(val(depth + 1) for _ in range(g.integers(0, 4))) + "]"
        if t == 8 and depth < 3:
            return "{" + ",".join('"k%d":%s' % (i, val(depth + 1)) for i in range(g.integers(0, 3))) + "}"
        return '"true"' if g.random() < 0.5 else '"false"'

    recs = []
    for _ in range(n):
        ks = KEYS + [LABEL, "extra1", "coolantTemp", "x"]
        g.shuffle(ks)
        ks = ks[:g.integers(0, len(ks) + 1)]
        if g.random() < 0.1 and ks:
            ks.append(ks[0])
        ws = lambda: " " * int(g.integers(0, 2))  # noqa: E731
        body = ",".join(ws() + '"%s"' % k + ws() + ":" + ws() + val() + ws() for k in ks)
        recs.append((ws() + "{" + body + "}" + ws()).encode())
    return recs, g
