#!/usr/bin/env python3
"""Deterministic synthetic-population generator for the covid-abm daily transmission cycle.

No population ships with the reference (its .gitignore excludes *.txt; pop/ holds only R scripts),
so every configuration in BASELINE.json is run on populations produced here. Two outputs:

* a dict of numpy arrays (structure-of-arrays + CSR) that feeds the C-ABI loader directly, and
* optionally the five text files in exactly the formats the reference parses
  (locations: Community.cpp:429-488, population: :200, comorbidity: :255, network: :513,
   public-activity: :285-296), so the unmodified reference runs on the same population.

Distributions are ours and are documented in BASELINE.md / SURVEY.md 8(d):
household size ~ 1+Poisson(1.5) truncated to 7; ages from a 9-bin US-like pyramid with an adult
first in every house; 5-17 year olds attend a school (sizes ~ lognormal, median 600); 60% of 18-64
year olds work (workplace sizes Zipf(2) capped at 5000, ~1% hospital staff, ~0.5% nursing-home
staff, school staff ~ 1 per 12 students); 0.4% of N live in nursing homes of ~100; houses are laid
out on a Morton curve, each house/NH maps to its nearest hospital, the household graph joins
houses that are close on the curve (mean degree ~4); 20% of workplaces take patrons (1/4 of those
high-risk), 40% of workplaces are non-essential; every non-LTC adult gets 5 patronised workplaces.
"""
from __future__ import annotations

import argparse
import os
import numpy as np

HOUSE, WORK, SCHOOL, HOSPITAL, NURSINGHOME = 0, 1, 2, 3, 4   # Parameters.h:74-81
PT_NONE, PT_LOW, PT_HIGH = 0, 1, 2                            # Parameters.h:98-103
TYPE_CODES = {HOUSE: "h", WORK: "w", SCHOOL: "s", HOSPITAL: "hf", NURSINGHOME: "n"}
N_PATRON = 5

_AGE_BINS = [(0, 4), (5, 17), (18, 29), (30, 44), (45, 54), (55, 64), (65, 74), (75, 84), (85, 100)]
_AGE_W = np.array([0.057, 0.163, 0.160, 0.195, 0.125, 0.130, 0.100, 0.050, 0.020])


def _morton(ix: np.ndarray, iy: np.ndarray) -> np.ndarray:
    def spread(v):
        v = v.astype(np.uint64) & np.uint64(0xFFFF)
        v = (v | (v << np.uint64(8))) & np.uint64(0x00FF00FF)
        v = (v | (v << np.uint64(4))) & np.uint64(0x0F0F0F0F)
        v = (v | (v << np.uint64(2))) & np.uint64(0x33333333)
        v = (v | (v << np.uint64(1))) & np.uint64(0x55555555)
        return v
    return spread(ix) | (spread(iy) << np.uint64(1))


def _draw_ages(rng, n, lo_bin=0, hi_bin=len(_AGE_BINS) - 1):
    w = _AGE_W[lo_bin:hi_bin + 1] / _AGE_W[lo_bin:hi_bin + 1].sum()
    b = rng.choice(np.arange(lo_bin, hi_bin + 1), size=n, p=w)
    lo = np.array([x[0] for x in _AGE_BINS])[b]
    hi = np.array([x[1] for x in _AGE_BINS])[b]
    return (lo + np.floor(rng.random(n) * (hi - lo + 1))).astype(np.int32)


def _csr_from_pairs(n_rows, rows, cols):
    order = np.lexsort((cols, rows))
    rows, cols = rows[order], cols[order]
    counts = np.bincount(rows, minlength=n_rows)
    off = np.zeros(n_rows + 1, dtype=np.int64)
    np.cumsum(counts, out=off[1:])
    return off, cols


def generate(n_target: int, seed: int = 20200210) -> dict:
    """Return the population as a dict of numpy arrays (see keys at the bottom)."""
    rng = np.random.default_rng(seed)
    n_nh_res = int(round(0.004 * n_target)) if n_target >= 5000 else 0
    n_nh = max(1, n_target // 30000) if n_nh_res > 0 else 0
    n_house_people = n_target - n_nh_res

    # ---- households -------------------------------------------------------------------------
    est = int(n_house_people / 2.4) + 16
    hh_size = np.minimum(1 + rng.poisson(1.5, size=est), 7)
    csum = np.cumsum(hh_size)
    n_house = int(np.searchsorted(csum, n_house_people, side="left")) + 1
    hh_size = hh_size[:n_house].copy()
    hh_size[-1] -= int(csum[n_house - 1] - n_house_people)
    assert hh_size.sum() == n_house_people and hh_size.min() >= 1

    # houses laid out along a Morton curve so that ids are spatially coherent
    side = 1 << 12
    hx = rng.integers(0, side, size=n_house); hy = rng.integers(0, side, size=n_house)
    order = np.argsort(_morton(hx, hy), kind="stable")
    hx, hy = hx[order], hy[order]
    scale = 2.0 / side   # coordinates in a 2x2-degree box
    house_x = -82.0 + hx * scale + rng.random(n_house) * scale
    house_y = 27.0 + hy * scale + rng.random(n_house) * scale

    home_h = np.repeat(np.arange(n_house, dtype=np.int64), hh_size)
    first_in_house = np.zeros(n_house_people, dtype=bool)
    first_in_house[np.concatenate(([0], np.cumsum(hh_size)[:-1]))] = True
    age_h = _draw_ages(rng, n_house_people)
    age_h[first_in_house] = _draw_ages(rng, int(first_in_house.sum()), lo_bin=2)

    # ---- nursing homes ------------------------------------------------------------------------
    nh_of_res = np.sort(rng.integers(0, max(n_nh, 1), size=n_nh_res)) if n_nh_res else np.zeros(0, dtype=np.int64)
    age_nh = _draw_ages(rng, n_nh_res, lo_bin=6)

    # location ids: houses, nursing homes, workplaces, schools, hospitals
    n_people = n_house_people + n_nh_res
    age = np.concatenate([age_h, age_nh]).astype(np.int32)
    home = np.concatenate([home_h, n_house + nh_of_res]).astype(np.int64)
    sex = (1 + (rng.random(n_people) < 0.51)).astype(np.int8)     # 1 male, 2 female (IPUMS)
    ltc = np.zeros(n_people, dtype=bool); ltc[n_house_people:] = True

    # ---- day locations ------------------------------------------------------------------------
    day = np.full(n_people, -1, dtype=np.int64)
    students = np.flatnonzero((age >= 5) & (age <= 17) & ~ltc)
    adults = np.flatnonzero((age >= 18) & (age <= 64) & ~ltc)
    employed = adults[rng.random(adults.size) < 0.60]
    rng.shuffle(employed)

    # schools: consecutive chunks of (Morton-ordered) students
    sizes = []
    tot = 0
    while tot < students.size:
        s = int(np.clip(rng.lognormal(np.log(600.0), 0.5), 50, 4000)); sizes.append(s); tot += s
    school_sizes = np.array(sizes, dtype=np.int64) if sizes else np.zeros(0, dtype=np.int64)
    n_school = school_sizes.size
    if n_school:
        school_sizes[-1] -= tot - students.size
        if school_sizes[-1] <= 0:
            school_sizes = school_sizes[:-1]; n_school -= 1
            school_sizes[-1] += students.size - school_sizes.sum()
    school_of_student = np.repeat(np.arange(n_school), school_sizes)

    n_hosp = max(1, n_target // 70000)
    n_school_staff = min(employed.size // 4, int(students.size / 12))
    n_hosp_staff = max(n_hosp * 2, int(0.01 * employed.size)) if employed.size > 10 else 0
    n_nh_staff = max(n_nh * 2, int(0.005 * employed.size)) if n_nh else 0
    n_special = n_school_staff + n_hosp_staff + n_nh_staff
    n_workers = employed.size - n_special
    assert n_workers >= 0

    # workplaces: Zipf(2) sizes capped at 5000
    wsizes = []
    tot = 0
    while tot < n_workers:
        chunk = np.minimum(rng.zipf(2.0, size=4096), 5000)
        wsizes.append(chunk); tot += int(chunk.sum())
    wsizes = np.concatenate(wsizes) if wsizes else np.zeros(0, dtype=np.int64)
    cs = np.cumsum(wsizes)
    n_work = int(np.searchsorted(cs, n_workers, side="left")) + 1 if n_workers > 0 else 0
    wsizes = wsizes[:n_work].astype(np.int64)
    if n_work:
        wsizes[-1] -= int(cs[n_work - 1] - n_workers)

    base_nh = n_house
    base_work = base_nh + n_nh
    base_school = base_work + n_work
    base_hosp = base_school + n_school
    n_loc = base_hosp + n_hosp

    day[students] = base_school + school_of_student
    p = 0
    day[employed[p:p + n_workers]] = base_work + np.repeat(np.arange(n_work), wsizes); p += n_workers
    if n_school_staff:
        w = school_sizes / school_sizes.sum()
        day[employed[p:p + n_school_staff]] = base_school + rng.choice(n_school, size=n_school_staff, p=w)
    p += n_school_staff
    if n_hosp_staff:
        day[employed[p:p + n_hosp_staff]] = base_hosp + (np.arange(n_hosp_staff) % n_hosp)
    p += n_hosp_staff
    if n_nh_staff:
        day[employed[p:p + n_nh_staff]] = base_nh + (np.arange(n_nh_staff) % n_nh)

    # ---- locations ----------------------------------------------------------------------------
    loc_type = np.empty(n_loc, dtype=np.uint8)
    loc_type[:n_house] = HOUSE; loc_type[base_nh:base_work] = NURSINGHOME
    loc_type[base_work:base_school] = WORK; loc_type[base_school:base_hosp] = SCHOOL
    loc_type[base_hosp:] = HOSPITAL
    loc_x = np.empty(n_loc); loc_y = np.empty(n_loc)
    loc_x[:n_house] = house_x; loc_y[:n_house] = house_y
    n_other = n_loc - n_house
    loc_x[n_house:] = -82.0 + rng.random(n_other) * 2.0
    loc_y[n_house:] = 27.0 + rng.random(n_other) * 2.0

    # compliance written with 6 decimals so that text and binary loaders see the same double
    comp_k = rng.integers(0, 1000001, size=n_loc)
    compliance = comp_k / 1e6
    compliance[n_house:] = -1.0
    risk = np.zeros(n_loc, dtype=np.float32)
    risk[:n_house] = (1.0 - compliance[:n_house]).astype(np.float32)      # Community.cpp:468, Location.h:85

    essential = np.ones(n_loc, dtype=np.uint8)
    essential[base_work:base_school] = (rng.random(n_work) >= 0.40)
    pubrisk = np.zeros(n_loc, dtype=np.uint8)
    public = rng.random(n_work) < 0.20
    if n_work and not public.any():
        public[int(np.argmax(wsizes))] = True
    high = public & (rng.random(n_work) < 0.25)
    pubrisk[base_work:base_school][public] = PT_LOW
    pubrisk[base_work:base_school][high] = PT_HIGH

    # nearest hospital for houses and nursing homes
    hosp_xy = np.stack([loc_x[base_hosp:], loc_y[base_hosp:]], axis=1)
    hospital = np.full(n_loc, -1, dtype=np.int64)
    res_locs = np.arange(base_work)
    try:
        from scipy.spatial import cKDTree
        _, nearest = cKDTree(hosp_xy).query(np.stack([loc_x[res_locs], loc_y[res_locs]], axis=1))
    except Exception:  # pragma: no cover - scipy is present in the image
        d = (loc_x[res_locs, None] - hosp_xy[None, :, 0]) ** 2 + (loc_y[res_locs, None] - hosp_xy[None, :, 1]) ** 2
        nearest = d.argmin(axis=1)
    hospital[res_locs] = base_hosp + nearest

    # household graph: join houses close on the Morton curve
    deg_half = rng.poisson(2.0, size=n_house)
    src = np.repeat(np.arange(n_house, dtype=np.int64), deg_half)
    off = rng.integers(1, 65, size=src.size)
    dst = src + off
    keep = dst < n_house
    e1, e2 = src[keep], dst[keep]
    pairs = np.unique(np.stack([e1, e2], axis=1), axis=0) if e1.size else np.zeros((0, 2), dtype=np.int64)
    e1, e2 = pairs[:, 0], pairs[:, 1]
    nbr_off, nbr = _csr_from_pairs(n_loc, np.concatenate([e1, e2]), np.concatenate([e2, e1]))

    # ---- patronised workplaces ----------------------------------------------------------------
    public_ids = base_work + np.flatnonzero(public)
    patron_people = np.flatnonzero((age >= 18) & ~ltc)                     # Community.cpp:288
    patron = np.full((n_people, N_PATRON), -1, dtype=np.int64)
    if public_ids.size:
        w = np.sqrt(wsizes[public]); w = w / w.sum()
        patron[patron_people] = public_ids[rng.choice(public_ids.size, size=(patron_people.size, N_PATRON), p=w)]

    comorbid_p = np.interp(age, [0, 100], [0.05, 0.45])
    comorbid = (rng.random(n_people) < comorbid_p).astype(np.uint8)

    # rosters (residents + workers) as CSR in the order the reference builds them:
    # loadPopulation pushes person p to home then to day location, in pid order (Community.cpp:219-224)
    has_day = day >= 0
    r_rows = np.concatenate([home, day[has_day]])
    r_cols = np.concatenate([np.arange(n_people, dtype=np.int64), np.flatnonzero(has_day)])
    roster_off, roster = _csr_from_pairs(n_loc, r_rows, r_cols)

    return dict(
        n_people=np.int64(n_people), n_loc=np.int64(n_loc),
        home=home.astype(np.int32), day=day.astype(np.int32), age=age.astype(np.uint8), sex=sex.astype(np.uint8),
        comorbid=comorbid, ltc=ltc.astype(np.uint8), patron=patron.astype(np.int32),
        loc_type=loc_type, loc_essential=essential, loc_pubrisk=pubrisk, loc_risk=risk,
        loc_hospital=hospital.astype(np.int32), loc_x=loc_x, loc_y=loc_y, loc_compliance=compliance,
        edges=np.stack([e1, e2], axis=1).astype(np.int32),
        nbr_off=nbr_off.astype(np.int64), nbr=nbr.astype(np.int32),
        roster_off=roster_off.astype(np.int64), roster=roster.astype(np.int32),
    )


def write_reference_files(pop: dict, out_dir: str, tag: str) -> dict:
    """Write the five text files the reference's loaders parse; returns their paths."""
    import pandas as pd
    os.makedirs(out_dir, exist_ok=True)
    paths = {k: os.path.join(out_dir, f"{k}-{tag}.txt") for k in
             ("locations", "population", "comorbidity", "network", "public-activity")}
    n_loc, n = int(pop["n_loc"]), int(pop["n_people"])
    tcode = np.array([TYPE_CODES[t] for t in range(5)], dtype=object)[pop["loc_type"]]
    ess = np.where(pop["loc_essential"] == 1, "y", "n").astype(object)
    ess[pop["loc_type"] != WORK] = "NA"
    pt = np.array(["N", "L", "H"], dtype=object)[pop["loc_pubrisk"]]
    comp = np.where(pop["loc_compliance"] >= 0, np.char.mod("%.6f", pop["loc_compliance"]), "-1")
    pd.DataFrame({"locid": np.arange(n_loc), "x": np.char.mod("%.6f", pop["loc_x"]), "y": np.char.mod("%.6f", pop["loc_y"]),
                  "type": tcode, "essential": ess, "hfid": pop["loc_hospital"], "compliance": comp,
                  "transmission": pt}).to_csv(paths["locations"], sep=" ", index=False)
    pd.DataFrame({"pid": np.arange(n), "hid": pop["home"], "sex": pop["sex"], "age": pop["age"],
                  "did": pop["day"]}).to_csv(paths["population"], sep=" ", index=False)
    pd.DataFrame({"pid": np.arange(n), "sex": pop["sex"], "age": pop["age"],
                  "undlycond": pop["comorbid"]}).to_csv(paths["comorbidity"], sep=" ", index=False)
    pd.DataFrame(pop["edges"]).to_csv(paths["network"], sep=" ", index=False, header=False)
    has = pop["patron"][:, 0] >= 0
    df = pd.DataFrame(pop["patron"][has]); df.insert(0, "pid", np.flatnonzero(has))
    df.to_csv(paths["public-activity"], sep=" ", index=False,
              header=["pid"] + [f"dest_locid_{i + 1}" for i in range(N_PATRON)])
    return paths


def main():
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--n", type=int, required=True, help="target number of agents")
    ap.add_argument("--seed", type=int, default=20200210)
    ap.add_argument("--out", required=True, help="output directory")
    ap.add_argument("--tag", default=None)
    ap.add_argument("--no-text", action="store_true", help="only write the .npz arrays")
    a = ap.parse_args()
    tag = a.tag or f"synth{a.n}"
    pop = generate(a.n, a.seed)
    os.makedirs(a.out, exist_ok=True)
    np.savez(os.path.join(a.out, f"pop-{tag}.npz"), **pop)
    if not a.no_text:
        write_reference_files(pop, a.out, tag)
    print(f"{tag}: people={int(pop['n_people'])} locations={int(pop['n_loc'])} "
          f"edges={pop['edges'].shape[0]} roster={pop['roster'].size}")


if __name__ == "__main__":
    main()
