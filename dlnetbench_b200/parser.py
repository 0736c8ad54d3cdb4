"""Parser for the proxies' stdout sections (the output-format contract).

Returns, per section id, the shapes the reference's plots/parser.py reads from ccutils'
MPIOutputParser: `json_data` (the global object) and `rank_outputs` (rank -> per-rank
object, what it finds under mpi_all_prints["ccutils_rank_json"])."""
from __future__ import annotations

import json


def parse_sections(text: str) -> dict:
    out: dict = {}
    for line in text.splitlines():
        if line.startswith("#CCUTILS_SECTION_BEGIN "):
            _, sid, title = line.split(" ", 2)
            out.setdefault(sid, {"title": title.strip('"'), "json_data": {}, "rank_outputs": {}})
        elif line.startswith("#CCUTILS_GLOBAL_JSON "):
            _, sid, js = line.split(" ", 2)
            out.setdefault(sid, {"title": "", "json_data": {}, "rank_outputs": {}})["json_data"] = json.loads(js)
        elif line.startswith("#CCUTILS_RANK_JSON "):
            _, sid, rank, js = line.split(" ", 3)
            out.setdefault(sid, {"title": "", "json_data": {}, "rank_outputs": {}})["rank_outputs"][int(rank)] = json.loads(js)
    return out
