"""CPU tests for row N2 (quota-table producer): libbgc's C++ restatement of the reference
synchronizer's sheet logic (src/synchronizer.rs:63-169, 208-281) against the independent Python one."""
import numpy as np
import pytest

from oracle import sheet_ref as S

HEADER = ["타임스탬프", "이메일 주소", "이름", "소속", "SNUCSE ID (id.snucse.org)", "사용할 서버를 선택해주세요", "필요한 GPU 개수",
          "필요한 vCPU 개수", "필요한 메모리 (GiB)", "필요한 스토리지 (GiB)", "필요한 MiG 개수", "요청 사유", "승인 여부"]


def _csv(rows, header=HEADER, nl="\n"):
    def cell(x):
        x = str(x)
        return '"' + x.replace('"', '""') + '"' if any(c in x for c in ',"\n\r') else x
    return nl.join(",".join(cell(c) for c in r) for r in [header] + rows) + nl


ROWS = [
    ["2023/08/01 1:00:00", "a@snu.ac.kr", "김철수", "컴퓨터공학부", "alice", "ferrari", 2, 16, 64, 500, 0, "연구, 실험", "O"],
    ["2023/08/01 2:00:00", "b@snu.ac.kr", "이영희", "전기정보", "bob", "ferrari, lamborghini", 1, 8, 32, 100, 1, 'say "hi"', " o "],
    ["2023/08/02 1:00:00", "c@snu.ac.kr", "박민수", "수리과학부", "carol", "lamborghini", 4, 32, 128, 500, 2, "multi\nline reason", "o"],
    ["2023/08/02 2:00:00", "a@snu.ac.kr", "김철수", "컴퓨터공학부", "alice", "ferrari", 8, 64, 256, 500, 0, "more", "X"],
    ["2023/08/03 1:00:00", "a@snu.ac.kr", "김철수", "컴퓨터공학부", "alice", "ferrari", 4, 32, 128, 100, 1, "update", "　O "],
    ["2023/08/03 2:00:00", "d@snu.ac.kr", "최지우", "컴퓨터공학부", "dave", "ferrari", "two", 8, 32, 100, 0, "bad int", "O"],
    ["2023/08/03 3:00:00", "e@snu.ac.kr", "정하늘", "컴퓨터공학부", "erin", "ferrari", "", 8, 32, 100, 0, "empty int", "O"],
    ["2023/08/03 4:00:00", "f@snu.ac.kr", "한가람", "컴퓨터공학부", "frank", "ferrari", -1, "+8", 32, 100, 0, "signs", "o"],
    ["2023/08/03 5:00:00", "g@snu.ac.kr", "오세훈", "컴퓨터공학부", "Alice", "ferrari", 1, 1, 1, 1, 0, "case differs", "O"],
    ["2023/08/04 1:00:00", "h@snu.ac.kr", "short row"],
    ["2023/08/04 2:00:00", "i@snu.ac.kr", "남궁민", "컴퓨터공학부", "ivan", "ferrari", 1, 4, 16, 100, 0, "spaces in int", " 1"],
]
ROWS[10][6] = " 1"


def test_header_inference_matches_reference_rules(bgc):
    for h in HEADER + ["SNUCSE ID", "GPU 개수", "vCPU 개수", "메모리", "이메일", "승인", "이름 ", "unknown", "", "MiG 개수와 GPU 개수"]:
        try:
            want = S.try_infer_header(h)
        except S.HeaderError:
            want = None
        assert bgc.sheet.infer_header(h) == want, h
    # order matters: "GPU 개수" is tested before "vCPU 개수" and "MiG 개수" (synchronizer.rs:114-131)
    assert bgc.sheet.infer_header("필요한 vCPU 개수") == "cpu_request"
    assert bgc.sheet.infer_header("MiG 개수와 GPU 개수") == "gpu_request"


@pytest.mark.parametrize("nl", ["\n", "\r\n"])
@pytest.mark.parametrize("bom", [False, True])
def test_parse_match_quota_against_python_restatement(bgc, nl, bom):
    text = ("﻿" if bom else "") + _csv(ROWS, nl=nl)
    want_rows, want_skipped = S.parse_csv(text)
    sh = bgc.sheet.Sheet(text)
    assert len(sh) == len(want_rows) == 7 and sh.skipped == want_skipped == 4
    for i, w in enumerate(want_rows):
        got = sh.row(i)
        assert {k: got[k] for k in S.FIELDS} == w
    names = ["alice", "bob", "carol", "dave", "frank", "Alice", "nobody", "ivan"]
    for server in ("ferrari", "lamborghini", "rrar", "", "mclaren"):
        assert sh.match(server, names) == S.match(want_rows, server, names), server
    # alice: the last AUTHORISED row wins (the "X" row is skipped, the later full-width-space "O" row counts)
    idx = sh.match("ferrari", ["alice"])[0]
    strings, canon = sh.quota(idx)
    assert strings == S.row_quota(want_rows[idx]) and list(strings) == sorted(strings, key=str.encode)
    assert strings == {"limits.cpu": "32", "limits.memory": "128Gi", "requests.cpu": "32", "requests.memory": "128Gi",
                       "requests.nvidia.com/gpu": "4", "requests.nvidia.com/mig-1g.10gb": "1", "requests.storage": "100Gi"}
    # the canonical integers are what Kubernetes would read from those strings
    for k, v in strings.items():
        unit = 1 if k.endswith("cpu") else 2 if ("memory" in k or "storage" in k) else 0
        assert bgc.pods.parse_quantity(v, unit) == canon[k]
    # negative and explicitly signed integers survive (serde parses them); quantity text follows format!()
    fi = sh.match("ferrari", ["frank"])[0]
    assert sh.quota(fi)[0]["requests.nvidia.com/gpu"] == "-1" and sh.quota(fi)[0]["requests.cpu"] == "8"


def test_unknown_header_fails_the_whole_parse(bgc):
    text = _csv(ROWS[:2], header=HEADER[:-1] + ["mystery column"])
    with pytest.raises(S.HeaderError):
        S.parse_csv(text)
    with pytest.raises(ValueError, match="unknown header"):
        bgc.sheet.Sheet(text)


def test_missing_required_column_skips_every_row(bgc):
    hdr = [h for h in HEADER if "MiG" not in h]
    rows = [[c for j, c in enumerate(r) if j != 10] for r in ROWS[:3]]
    text = _csv(rows, header=hdr)
    assert S.parse_csv(text) == ([], 3)
    sh = bgc.sheet.Sheet(text)
    assert len(sh) == 0 and sh.skipped == 3


def test_hard_table_feeds_the_quota_engine_shapes(bgc):
    sh = bgc.sheet.Sheet(_csv(ROWS))
    keys = ["requests.nvidia.com/gpu", "requests.cpu", "requests.memory"]
    names = ["alice", "bob", "nobody", "carol"]
    hard, mask, idx = sh.hard_table("ferrari", names, keys)
    assert idx[2] == -1 and mask.tolist() == [7, 7, 0, 0]            # carol only asked for lamborghini
    assert hard[:, 0].tolist() == [4, 32_000, 128 << 30] and hard[:, 1].tolist() == [1, 8_000, 32 << 30]
    assert hard.dtype == np.int64 and hard.shape == (3, 4)


def test_large_sheet_random_against_python(bgc):
    rng = np.random.default_rng(0)
    users = [f"user{i}" for i in range(300)]
    rows = []
    for i in range(3000):
        auth = rng.choice(["O", "o", "X", "", " O", "o\t", "Ｏ", "0"])
        srv = rng.choice(["ferrari", "lamborghini", "ferrari, lamborghini", "bentley"])
        rows.append([f"t{i}", "m@x", "n", "d", rng.choice(users), srv, int(rng.integers(0, 9)), int(rng.integers(1, 65)),
                     int(rng.integers(1, 257)), int(rng.integers(1, 501)), int(rng.integers(0, 3)), "r", auth])
    text = _csv(rows)
    want_rows, _ = S.parse_csv(text)
    sh = bgc.sheet.Sheet(text)
    assert len(sh) == len(want_rows) == 3000
    for server in ("ferrari", "lamborghini"):
        got = sh.match(server, users)
        assert got == S.match(want_rows, server, users)
        for ri in got[:50]:
            if ri >= 0:
                assert sh.quota(ri)[0] == S.row_quota(want_rows[ri])
