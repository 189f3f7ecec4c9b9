"""Social-traffic wire format.  The SUMO pose conversion is pinned by golden
values produced by the reference's own Heading.from_sumo / Pose.from_front_bumper
(tests/golden/gen_golden.py, traffic.npz)."""
import os
import sqlite3

import numpy as np

from conftest import GOLDEN
from smarts_b200 import traffic


def test_sumo_conversion_matches_reference_golden():
    z = np.load(os.path.join(GOLDEN, "traffic.npz"))
    recs = traffic.records_from_sumo([str(i) for i in range(len(z["angle"]))], z["front"], z["angle"], z["speed"],
                                     [str(c) for c in z["vclass"]])
    got = np.array([[r.x, r.y, r.heading, r.length, r.width, r.height] for r in recs])
    assert np.allclose(got[:, :2], z["centre"], rtol=0, atol=1e-12)
    assert np.allclose(got[:, 2], z["heading"], rtol=0, atol=1e-15)
    assert np.array_equal(got[:, 3:], z["dims"])


def test_slot_assignment_is_stable_and_marks_departures():
    R = traffic.VehicleRecord
    mk = lambda vid, x: R(vid, x, 0.0, 0.0, 5.0, 3.68, 1.47, 1.4)  # noqa: E731
    ticks = [[[mk("a", 1), mk("b", 2)]], [[mk("b", 3), mk("c", 4)]], [[mk("c", 5), mk("a", 6), mk("d", 7), mk("e", 8)]]]
    rec = traffic.recording_from_ticks(ticks, M=3)
    assert rec.shape == (3, 1, 3, 8)
    assert rec[0, 0, :, 7].tolist() == [1, 1, 0] and rec[0, 0, 0, 0] == 1 and rec[0, 0, 1, 0] == 2
    # b keeps slot 1; a left: c takes a fresh slot, a's old slot stays dead this tick
    assert rec[1, 0, 1, 0] == 3 and rec[1, 0, 2, 0] == 4 and rec[1, 0, 0, 7] == 0
    # c keeps slot 2; a returns into the slot that cooled down; b's slot is dead this tick, so d and e do not fit
    assert rec[2, 0, 2, 0] == 5 and rec[2, 0, 0, 0] == 6 and rec[2, 0, :, 7].tolist() == [1, 0, 1]
    asg = traffic.SlotAssigner(3)
    for t in ticks:
        asg.tick(t[0])
    assert asg.dropped == 2


def test_history_sqlite_replay(tmp_path):
    p = str(tmp_path / "hist.db")
    con = sqlite3.connect(p)
    con.execute("CREATE TABLE Vehicle (id INTEGER PRIMARY KEY, type INTEGER NOT NULL, length REAL, width REAL, "
                "height REAL, is_ego_vehicle INTEGER DEFAULT 0) WITHOUT ROWID")
    con.execute("CREATE TABLE Trajectory (vehicle_id INTEGER NOT NULL, sim_time REAL NOT NULL, position_x REAL NOT NULL, "
                "position_y REAL NOT NULL, heading_rad REAL NOT NULL, speed REAL DEFAULT 0.0, lane_id INTEGER DEFAULT 0, "
                "PRIMARY KEY (vehicle_id, sim_time)) WITHOUT ROWID")
    con.execute("INSERT INTO Vehicle VALUES (7, 2, 4.5, NULL, -1, 0)")   # width/height fall back to passenger
    con.execute("INSERT INTO Vehicle VALUES (9, 3, NULL, NULL, NULL, 0)")  # truck defaults
    for k in range(1, 6):
        con.execute("INSERT INTO Trajectory VALUES (7, ?, ?, 1.0, 4.0, 10.0, 0)", (round(0.1 * k, 1), 10.0 * k))
    for k in range(3, 6):
        con.execute("INSERT INTO Trajectory VALUES (9, ?, ?, 2.0, -0.5, 3.0, 0)", (round(0.1 * k, 1), 5.0 * k))
    con.commit(); con.close()
    rec = traffic.load_history_sqlite(p, dt=0.1, n_ticks=6, M=4, n_envs=2, env_time_offset=0.2)
    assert rec.shape == (6, 2, 4, 8)
    assert rec[0, 0, :, 7].sum() == 0                        # (-0.1, 0.0]: nothing yet
    assert rec[1, 0, 0].tolist() == [10.0, 1.0, 4.0 - 2 * np.pi, 10.0, 4.5, 1.47, 1.4, 1.0]
    assert rec[3, 0, :, 7].sum() == 2 and rec[3, 0, 1, 4:7].tolist() == [5.0, 1.91, 1.89]
    assert np.array_equal(rec[2:, 0, 0], rec[:4, 1, 0])      # env 1 runs 0.2 s ahead: same vehicle, same slot
