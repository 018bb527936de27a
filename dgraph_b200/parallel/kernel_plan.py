"""Kernel-side graph structures derived from a :class:`CommunicationPattern`.

The sm_100a kernels do not consume COO edge lists: aggregation walks a **CSR** sorted by the owned
(central) vertex so every output row is reduced by one warp without atomics (deterministic), and the
backward pass walks the **CSC** (sorted by neighbour row over ``[local | halo]``) for the transposed
reduction.  Indices are int32 (per-rank shards stay below 2^31 rows/edges).

No counterpart in the reference, which keeps COO + ``scatter_add`` atomics (``experiments/OGB/GCN.py:57-65``).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch


MAX_COUNTED_DEGREE = 65535  # uint16 per-(row, column) active-edge counters of the add-ReLU aggregation kernel
# One lane group walks one CSR / CSC segment.  Segments longer than HUB_DEGREE (hub vertices of a power-law graph: ogbn-products has
# a vertex of degree 17,481) are cut into chunks of HUB_CHUNK slots that run as rows of their own in a second, small launch and are
# added up afterwards, so a hub costs HUB_CHUNK slots of serial work instead of its whole degree.
HUB_DEGREE = int(os.environ.get("DGRAPH_B200_HUB_DEGREE", "2048"))
HUB_CHUNK = int(os.environ.get("DGRAPH_B200_HUB_CHUNK", "512"))


@dataclass
class HubPlan:
    hubs: torch.Tensor  # int64 [n_hubs]  segments longer than HUB_DEGREE
    vrow: torch.Tensor  # int64 [n_virtual]  segment each chunk belongs to
    vbeg: torch.Tensor  # int32 [n_virtual]  first slot of the chunk
    vend: torch.Tensor  # int32 [n_virtual]  one past its last slot
    main_end: torch.Tensor  # int32 [n_segments]  end pointers of the main launch: hub segments are empty there
    seg: torch.Tensor  # int64 [n_virtual]  index into `hubs` of each chunk


@dataclass
class GraphKernelPlan:
    num_local: int  # rows that are aggregated into (central side)
    num_nbr_local: int  # local rows of the neighbour side (== num_local for homogeneous graphs)
    num_halo: int
    num_edges: int
    # CSR over central vertices
    rowptr: torch.Tensor  # int32 [num_local + 1]  (int64 when E >= 2^31)
    col: torch.Tensor  # int32 [E]   neighbour row in [nbr_local | halo] numbering
    csr_eid: torch.Tensor  # int32/64 [E] original edge id of each CSR slot
    # CSC over neighbour rows (local + halo)
    colptr: torch.Tensor  # [num_nbr_local + num_halo + 1]
    row: torch.Tensor  # int32 [E]   central row of each CSC slot
    csc_eid: torch.Tensor  # [E] original edge id of each CSC slot
    csc_to_csr: torch.Tensor  # [E] CSR slot of each CSC slot
    # halo addressing (pull): owner rank and row inside the owner's matrix
    halo_owner: Optional[torch.Tensor]  # int32 [num_halo]
    halo_remote_row: Optional[torch.Tensor]  # int32 [num_halo]
    # send side
    send_local_idx: torch.Tensor  # int32 [total_sends]

    @property
    def num_aug(self) -> int:
        return self.num_nbr_local + self.num_halo

    def max_row_degree(self) -> int:
        """Largest CSR segment (edges of one central vertex); one device-to-host read, cached.  The fused GraphConv aggregation
        counts active edges per (row, column) in 16 bits, so its callers check this against ``MAX_COUNTED_DEGREE``."""
        d = getattr(self, "_max_row_degree", None)
        if d is None:
            rp = self.rowptr
            d = int((rp[1:] - rp[:-1]).max()) if rp.numel() > 1 else 0
            self._max_row_degree = d
        return d

    def hub_plan(self, side: str = "csr") -> Optional[HubPlan]:
        """Chunks of the over-long segments of the CSR (``side="csr"``) or CSC walk, or None when there are none (one device-to-host
        read, cached)."""
        cache = self.__dict__.setdefault("_hub_plans", {})
        if side in cache:
            return cache[side]
        ptr = (self.rowptr if side == "csr" else self.colptr).long()
        deg = ptr[1:] - ptr[:-1]
        hubs = torch.nonzero(deg > HUB_DEGREE, as_tuple=True)[0]
        plan = None
        if hubs.numel():
            nch = (deg[hubs] + HUB_CHUNK - 1) // HUB_CHUNK
            seg = torch.repeat_interleave(torch.arange(hubs.numel(), device=ptr.device), nch)
            first = torch.cumsum(nch, 0) - nch
            k = torch.arange(int(nch.sum()), device=ptr.device) - first[seg]
            vrow = hubs[seg]
            vbeg = ptr[vrow] + k * HUB_CHUNK
            vend = torch.minimum(vbeg + HUB_CHUNK, ptr[vrow + 1])
            main_end = ptr[1:].clone()
            main_end[hubs] = ptr[hubs]
            it = self.rowptr.dtype
            plan = HubPlan(hubs, vrow, vbeg.to(it).contiguous(), vend.to(it).contiguous(), main_end.to(it).contiguous(), seg)
        cache[side] = plan
        return plan

    def csr_to_csc(self) -> torch.Tensor:
        """CSC slot of each CSR slot (inverse of ``csc_to_csr``), built on first use.  The forward add-ReLU kernel writes its
        ReLU bit masks at the CSC slot so that the backward CSC walk reads them as one sequential stream."""
        t = getattr(self, "_csr_to_csc", None)
        if t is None or t.device != self.csc_to_csr.device:
            t = torch.empty_like(self.csc_to_csr)
            t[self.csc_to_csr.long()] = torch.arange(self.csc_to_csr.numel(), device=t.device, dtype=t.dtype)
            self._csr_to_csc = t
        return t

    def col_peer(self, shift: int) -> torch.Tensor:
        """``col`` for the pull kernels: a halo neighbour ``h`` becomes ``~((owner[h] << shift) | row_on_owner[h])`` (negative), so
        the kernel forms the NVLink address from the index alone -- no table look-ups behind the index load.  Local neighbours
        keep their row number.  Cached per ``shift``."""
        cache = self.__dict__.setdefault("_col_peer", {})
        t = cache.get(shift)
        if t is None or t.device != self.col.device:
            assert self.halo_owner is not None and self.halo_remote_row is not None, "plan was built without halo addressing"
            nl = self.num_nbr_local
            dev = self.col.device
            code = ~((self.halo_owner.to(dev).long() << shift) | self.halo_remote_row.to(dev).long())
            table = torch.cat([torch.arange(nl, device=dev, dtype=torch.long), code])
            t = table[self.col.long()].to(torch.int32).contiguous()
            cache[shift] = t
        return t

    def remote_edge_count(self) -> int:
        """Edges whose neighbour row lives on another rank (each costs one NVLink row read in the pull kernels)."""
        c = getattr(self, "_remote_edges", None)
        if c is None:
            c = int((self.col >= self.num_nbr_local).sum())
            self._remote_edges = c
        return c

    def rowmid(self) -> torch.Tensor:
        """First halo-neighbour slot of every CSR row (rows are neighbour-sorted, local neighbours first): the local part of
        row r is ``[rowptr[r], rowmid[r])``, the halo part ``[rowmid[r], rowptr[r+1])``."""
        t = getattr(self, "_rowmid", None)
        if t is None or t.device != self.col.device:
            is_local = (self.col < self.num_nbr_local).to(torch.long)
            pref = torch.zeros(self.col.numel() + 1, dtype=torch.long, device=self.col.device)
            torch.cumsum(is_local, 0, out=pref[1:])
            rp = self.rowptr.long()
            t = (rp[:-1] + (pref[rp[1:]] - pref[rp[:-1]])).to(self.rowptr.dtype).contiguous()
            self._rowmid = t
        return t

    def restricted_csc(self, rows: torch.Tensor):
        """CSC structure restricted to the edges whose *row* (central vertex) is in ``rows``: ``(colptr, row, csc_to_csr)``.
        Used by the backward of the last message-passing layer when the incoming gradient is zero outside ``rows`` (the
        labelled vertices).  Built once per index tensor and cached."""
        # The cache keeps the index tensor alive: as long as an entry exists its memory cannot be recycled for another index
        # tensor, so "same address, same length, version unchanged since the entry was made" means "same contents".
        cache = self.__dict__.setdefault("_restricted", {})
        key = (rows.data_ptr(), rows.numel())
        hit = cache.get(key)
        if hit is not None and hit[0]._version == hit[1] == rows._version and hit[0].device == rows.device:
            return hit[2]
        dev = self.row.device
        member = torch.zeros(self.num_local + 1, dtype=torch.bool, device=dev)
        member[rows.to(dev).long()] = True
        keep = member[self.row.long()]
        pref = torch.zeros(self.row.numel() + 1, dtype=torch.long, device=dev)
        torch.cumsum(keep, 0, out=pref[1:])
        colptr = pref[self.colptr.long()].to(self.colptr.dtype)
        out = (colptr, self.row[keep].contiguous(), self.csc_to_csr[keep].contiguous())
        if len(cache) >= 4:
            cache.clear()
        cache[key] = (rows, rows._version, out)
        return out

    @staticmethod
    def from_edges(edge_list: torch.Tensor, num_local: int, num_aug: int, device=None, num_nbr_local=None,
                   halo_owner=None, halo_remote_row=None, send_local_idx=None) -> "GraphKernelPlan":
        dev = torch.device(device) if device is not None else edge_list.device
        e = edge_list.to(dev)
        E = int(e.shape[0])
        idx_t = torch.int32 if E < 2 ** 31 else torch.int64
        src = e[:, 0].to(torch.long)
        dst = e[:, 1].to(torch.long)
        # CSR: stable sort by (src, dst) gives neighbour-sorted rows => better L2 locality on gathers
        key = src * max(num_aug, 1) + dst
        csr_perm = torch.argsort(key, stable=True)
        rowptr = torch.zeros(num_local + 1, dtype=torch.long, device=dev)
        rowptr[1:] = torch.bincount(src, minlength=num_local).cumsum(0)
        col = dst[csr_perm]
        key2 = dst * max(num_local, 1) + src
        csc_perm = torch.argsort(key2, stable=True)
        colptr = torch.zeros(num_aug + 1, dtype=torch.long, device=dev)
        colptr[1:] = torch.bincount(dst, minlength=num_aug).cumsum(0)
        row = src[csc_perm]
        inv_csr = torch.empty(E, dtype=torch.long, device=dev)
        inv_csr[csr_perm] = torch.arange(E, device=dev)
        csc_to_csr = inv_csr[csc_perm]
        nl = num_local if num_nbr_local is None else num_nbr_local
        return GraphKernelPlan(
            num_local=num_local,
            num_nbr_local=nl,
            num_halo=num_aug - nl,
            num_edges=E,
            rowptr=rowptr.to(idx_t),
            col=col.to(torch.int32),
            csr_eid=csr_perm.to(idx_t),
            colptr=colptr.to(idx_t),
            row=row.to(torch.int32),
            csc_eid=csc_perm.to(idx_t),
            csc_to_csr=csc_to_csr.to(idx_t),
            halo_owner=None if halo_owner is None else halo_owner.to(dev, torch.int32),
            halo_remote_row=None if halo_remote_row is None else halo_remote_row.to(dev, torch.int32),
            send_local_idx=(torch.empty(0, dtype=torch.int32, device=dev) if send_local_idx is None
                            else send_local_idx.to(dev, torch.int32)),
        )

    @staticmethod
    def from_pattern(pattern, device=None) -> "GraphKernelPlan":
        nl = pattern.num_neighbor_local_vertices
        if nl is None:
            nl = pattern.num_local_vertices
        plan = GraphKernelPlan.from_edges(
            pattern.local_edge_list,
            pattern.num_local_vertices,
            nl + pattern.num_halo_vertices,
            device=device,
            num_nbr_local=nl,
            halo_owner=pattern.halo_owner,
            halo_remote_row=pattern.halo_remote_row,
            send_local_idx=pattern.send_local_idx,
        )
        from dgraph_b200.utils import debug

        if debug.validate_enabled():  # DGRAPH_B200_VALIDATE=1: range-check every index the kernels will dereference
            debug.validate_pattern(pattern)
            debug.validate_kernel_plan(plan)
        return plan
