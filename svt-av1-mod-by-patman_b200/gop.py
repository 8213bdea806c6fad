"""Host-side job planner: which pictures are ME'd against which references.

A compact model of the reference's random-access prediction structure as the PAME stage sees it
(Source/Lib/Codec/pd_process.c:3463-3549 posts one TASK_PAME per non-intra picture with
ref_list{0,1}_count_try references, me_process.c:219-262 resolves them to pa-ref pyramids):
hierarchical mini-GOPs of 2^levels pictures, coded layer by layer; list 0 = nearest already-coded
pictures in the past, list 1 = nearest in the future; base-layer pictures use past pictures in
both lists. It is a workload generator for tests / bench / sharding, not an encoder decision.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Tuple


@dataclass
class PictureJob:
    picture_number: int
    temporal_layer: int
    is_ref: bool
    refs: Dict[Tuple[int, int], int] = field(default_factory=dict)  # (list, idx) -> picture number

    @property
    def n_l0(self):
        return 1 + max(ri for (li, ri) in self.refs if li == 0)

    @property
    def n_l1(self):
        return 1 + max([ri for (li, ri) in self.refs if li == 1], default=-1)


def _nearest(avail, x, n, past):
    cand = sorted([p for p in avail if (p < x if past else p > x)], key=lambda p: abs(p - x))
    return cand[:n]


def ra_jobs(n_frames: int, levels: int = 5, n_l0: int = 3, n_l1: int = 2) -> List[PictureJob]:
    """Jobs in decode order for pictures 1..n_frames-1 (picture 0 is intra: no ME)."""
    jobs: List[PictureJob] = []
    coded = [0]
    base = 0
    while base < n_frames - 1:
        remaining = n_frames - 1 - base
        lv = levels
        while (1 << lv) > remaining:
            lv -= 1
        L = 1 << lv
        order = [(L, 0)]  # (position, layer)
        for t in range(1, lv + 1):
            step = L >> t
            order += [(k, t) for k in range(step, L, 2 * step)]
        for pos, t in order:
            x = base + pos
            tl = t + (levels - lv) if t > 0 else 0
            past = _nearest(coded, x, n_l0, True)
            while len(past) < n_l0:
                past.append(past[-1])
            refs = {(0, i): p for i, p in enumerate(past)}
            if n_l1:
                if t == 0:
                    fut = list(past[:n_l1])  # base layer: list 1 mirrors the past references
                else:
                    fut = _nearest(coded, x, n_l1, False)
                    fill = [p for p in past if p not in fut]
                    while len(fut) < n_l1:
                        fut.append(fill.pop(0) if fill else past[0])
                refs.update({(1, i): p for i, p in enumerate(fut)})
            jobs.append(PictureJob(x, tl, tl < levels, refs))
            coded.append(x)
        base += L
    return jobs


def shard_jobs(jobs: List[PictureJob], world_size: int, rank: int) -> List[PictureJob]:
    """Picture sharding across GPUs: job i (decode order) -> rank i % world_size."""
    return [j for i, j in enumerate(jobs) if i % world_size == rank]


def pictures_needed(jobs: List[PictureJob]) -> List[int]:
    need = set()
    for j in jobs:
        need.add(j.picture_number)
        need.update(j.refs.values())
    return sorted(need)
