"""Workload shapes of BASELINE.json.configs and synthetic-data generators.

Table-size lists: Criteo-Kaggle from the reference's run scripts
(examples/cpp/DLRM/run_dlrm_kaggle_day1.sh:5); MLPerf DLRM-Terabyte (40M cap)
from the public MLPerf config (not in the reference tree).
"""
import math

KAGGLE_ROWS = [1461, 584, 10131227, 2202608, 306, 24, 12518, 634, 4, 93146, 5684, 8351593, 3195, 28,
               14993, 5461306, 11, 5653, 2173, 4, 7046547, 18, 16, 286181, 105, 142572]

TERABYTE_ROWS = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346,
                 10, 2208, 11938, 155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108,
                 36]

AGGR_SUM, AGGR_AVG = 21, 22

WORKLOADS = {
    # name: rows per table, out_dim, bag size, aggr, batch
    "C1": dict(rows=[100000] * 8, dim=64, bag=1, aggr=AGGR_SUM, batch=256,
               desc="DLRM example small: 8 tables x 100k x 64, bag 1, SUM, batch 256"),
    "C2": dict(rows=KAGGLE_ROWS, dim=16, bag=1, aggr=AGGR_SUM, batch=2048,
               desc="Criteo-Kaggle shape: 26 tables (max 10.1M rows) x 16, one-hot, batch 2048"),
    "C3": dict(rows=TERABYTE_ROWS, dim=128, bag=1, aggr=AGGR_SUM, batch=65536,
               desc="MLPerf DLRM-Terabyte shape: 26 tables (max 40.0M rows) x 128, one-hot, batch 65536"),
    "C4": dict(rows=[10000000] * 8, dim=128, bag=32, aggr=AGGR_AVG, batch=16384,
               desc="multi-hot stress: 8 tables x 10M x 128, pooling 32, AVG, batch 16384"),
    "C5": dict(rows=[1000000] * 64, dim=64, bag=1, aggr=AGGR_SUM, batch=65536,
               desc="Zipf sweep shape: 64 tables x 1M x 64, one-hot, batch 65536 (batch is our choice)"),
}


def table_bytes(w):
    return sum(w["rows"]) * w["dim"] * 4


def lookups_per_step(w, tables=None):
    n = len(w["rows"]) if tables is None else len(tables)
    return n * w["batch"] * w["bag"]


def gen_indices(torch, rows, batch, bag, dist="uniform", alpha=0.0, generator=None, device="cuda"):
    """[batch, bag] int64 rows in [0, rows).  dist 'uniform' (the reference
    draws std::rand() % N, dlrm.cc:419-420) or 'zipf': rank r in 1..rows with
    p(r) ~ r^-alpha by inverse transform of the continuous bounded power law,
    rank -> row by a fixed multiplicative bijection so hot rows are scattered."""
    if dist == "uniform" or alpha == 0.0:
        return torch.randint(0, rows, (batch, bag), dtype=torch.int64, device=device, generator=generator)
    u = torch.rand((batch, bag), dtype=torch.float64, device=device, generator=generator)
    if abs(alpha - 1.0) < 1e-9:
        rank = torch.exp(u * math.log(rows + 1.0))
    else:
        a = 1.0 - alpha
        rank = ((float(rows + 1) ** a - 1.0) * u + 1.0) ** (1.0 / a)
    rank = rank.floor().to(torch.int64).clamp_(1, rows) - 1
    mult = _coprime_multiplier(rows)
    return (rank * mult + 12345) % rows


def _coprime_multiplier(n):
    m = max(3, int(n * 0.6180339887) | 1)
    while math.gcd(m, n) != 1:
        m += 2
    return m
