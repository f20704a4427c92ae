"""oracle/sheet_ref.py — TEST INFRASTRUCTURE ONLY.  Restates the synchronizer's sheet logic,
/root/reference/src/synchronizer.rs:63-169 (Row, try_infer_header, parse_csv) and :208-281 (server
filter, last authorised matching row, the seven `hard` quantities), with Python's own csv module and
str methods so that it shares nothing with the C++ in libbgc (csrc/bgc_host.cpp).

PARITY UNPINNED: the reference has no tests or fixtures for this path and cannot be built here; the
rust `csv` 1.2 reader and serde's integer parsing are restated from their documentation.
"""
from __future__ import annotations

import csv
import io
import re

FIELDS = ("name", "department", "id_username", "gpu_server", "gpu_request", "cpu_request", "memory_request",
          "storage_request", "mig_request", "authorized")
INT_FIELDS = ("gpu_request", "cpu_request", "memory_request", "storage_request", "mig_request")


class HeaderError(ValueError):
    pass


def try_infer_header(header: str) -> str:
    """synchronizer.rs:97-143 — first match wins."""
    if header == "타임스탬프":
        return "timestamp"
    if header == "이름":
        return "name"
    if header == "소속":
        return "department"
    for needle, name in (("SNUCSE ID", "id_username"), ("사용할 서버", "gpu_server"), ("GPU 개수", "gpu_request"),
                         ("vCPU 개수", "cpu_request"), ("메모리", "memory_request"), ("스토리지", "storage_request"),
                         ("MiG 개수", "mig_request"), ("요청 사유", "description"), ("승인", "authorized"), ("이메일", "email")):
        if needle in header:
            return name
    raise HeaderError(f'unknown header: "{header}"')


_INT = re.compile(r"^[+-]?[0-9]+$")


def parse_csv(content: str) -> tuple:
    """synchronizer.rs:145-169.  Returns (rows, skipped)."""
    if content.startswith("﻿"):
        content = content[1:]
    records = [r for r in csv.reader(io.StringIO(content, newline="")) if r]
    if not records:
        return [], 0
    names = [try_infer_header(h) for h in records[0]]
    rows, skipped = [], 0
    for rec in records[1:]:
        if len(rec) != len(names) or any(names.count(f) != 1 for f in FIELDS):
            skipped += 1
            continue
        d = dict(zip(names, rec))
        try:
            for f in INT_FIELDS:
                if not _INT.match(d[f]) or not -(1 << 63) <= int(d[f]) < (1 << 63):
                    raise ValueError(f)
                d[f] = int(d[f])
        except ValueError:
            skipped += 1  # "row parsing error. skipping" :160-163
            continue
        rows.append({f: d[f] for f in FIELDS})
    return rows, skipped


def match(rows: list, server_name: str, names: list) -> list:
    """synchronizer.rs:208-236: index (into `rows`) of the last authorised row of this server whose
    id_username equals the name, or -1."""
    mine = [i for i, r in enumerate(rows) if server_name in r["gpu_server"]]
    out = []
    for name in names:
        hit = -1
        for i in reversed(mine):
            if rows[i]["authorized"].strip().lower() == "o" and rows[i]["id_username"] == name:
                hit = i
                break
        out.append(hit)
    return out


def row_quota(row: dict) -> dict:
    """synchronizer.rs:249-281: the BTreeMap of quantity strings (sorted keys)."""
    q = {
        "requests.cpu": str(row["cpu_request"]),
        "requests.memory": f'{row["memory_request"]}Gi',
        "limits.cpu": str(row["cpu_request"]),
        "limits.memory": f'{row["memory_request"]}Gi',
        "requests.nvidia.com/gpu": str(row["gpu_request"]),
        "requests.storage": f'{row["storage_request"]}Gi',
        "requests.nvidia.com/mig-1g.10gb": str(row["mig_request"]),
    }
    return dict(sorted(q.items(), key=lambda kv: kv[0].encode()))
