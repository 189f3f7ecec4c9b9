"""The reference's own acceptance test for the grid maps, /root/reference/smarts/core/tests/test_observations.py:93-233,
re-expressed on the CUDA path through the hiway-v0 surface: scenarios/loop, one Laner agent (keep_lane) stepping 30
ticks among social traffic, then
  * the ego and every neighbour project onto occupied OGM pixels and onto drivable-area pixels (tolerance 2 px,
    `sample_vehicle_pos`, :136-155);
  * the first and the last waypoint of every road-waypoint path project onto drivable-area pixels (:198-233).
`project_2d` (:93-126) is restated with the orthographic lens written out (Panda3D is unavailable on this box: the lens
maps film coordinates x / (film_w / 2), z / (film_h / 2) to [-1, 1]).  Differences from the reference's sizes, both
caps of this library: the grids are 1024 x 1024 (the reference test uses 1536 x 1536) at the same 50 / 256 m per pixel,
and the road-waypoint horizon is 32 (the reference test uses 40); the RGB camera needs the renderer and is left out.
The grids themselves stay "parity unpinned" (DESIGN.md 2): this is the reference's projection contract, not its pixels."""
import numpy as np
import pytest

from conftest import load_map
from smarts_b200 import scenario_gen
from smarts_b200.agent_interface import (OGM, ActionSpaceType, Agent, AgentInterface, AgentSpec, DoneCriteria,
                                         DrivableAreaGridMap, NeighborhoodVehicles, RoadWaypoints)
from smarts_b200.hiway_env import HiWayEnv, Scenario
from smarts_b200.observation import Mission

pytestmark = pytest.mark.gpu

AGENT_ID = "Agent-007"
NUM_STEPS = 30
MAP_WIDTH = MAP_HEIGHT = 1024
HALF_WIDTH, HALF_HEIGHT = MAP_WIDTH / 2, MAP_HEIGHT / 2
MAP_RESOLUTION = 50 / 256


def project_2d(img_metadata, pos):
    center = np.array(img_metadata.camera_pos)
    heading = np.radians(img_metadata.camera_heading_in_degrees)
    p = np.asarray(pos, float) - center
    xr = p[0] * np.cos(-heading) - p[1] * np.sin(-heading)
    yr = p[0] * np.sin(-heading) + p[1] * np.cos(-heading)
    # OrthographicLens.project with film size (res * W, res * H): inside the frustum iff both are within [-1, 1]
    nx, ny = xr / (0.5 * MAP_RESOLUTION * MAP_WIDTH), yr / (0.5 * MAP_RESOLUTION * MAP_HEIGHT)
    x, y = int(HALF_HEIGHT), int(HALF_WIDTH)
    inside = abs(nx) <= 1 and abs(ny) <= 1
    if inside:
        x = int(-ny * HALF_HEIGHT + HALF_HEIGHT)
        y = int(nx * HALF_WIDTH + HALF_WIDTH)
    return x, y, inside


def apply_tolerance(arr, x, y, tolerance):
    return arr[max(x - tolerance, 0):x + tolerance, max(y - tolerance, 0):y + tolerance, :]


def test_observations():
    rm = load_map("loop")
    itf = AgentInterface(road_waypoints=RoadWaypoints(32),
                         neighborhood_vehicles=NeighborhoodVehicles(radius=max(MAP_WIDTH, MAP_HEIGHT) * MAP_RESOLUTION * 0.5),
                         drivable_area_grid_map=DrivableAreaGridMap(width=MAP_WIDTH, height=MAP_HEIGHT, resolution=MAP_RESOLUTION),
                         ogm=OGM(width=MAP_WIDTH, height=MAP_HEIGHT, resolution=MAP_RESOLUTION),
                         # the injected traffic does not yield: a contact must not end the episode before tick 30
                         done_criteria=DoneCriteria(collision=False, off_road=False, off_route=False),
                         action=ActionSpaceType.Lane)
    spec = AgentSpec(interface=itf, agent_builder=lambda: Agent.from_function(lambda _: "keep_lane"))
    traffic = scenario_gen.SocialTraffic(rm, 1, 40, 42)
    sc = Scenario(rm, {AGENT_ID: Mission("445633931", 1, 30.0, "445633932", start_speed=8.0)}, traffic, n_social=40)
    env = HiWayEnv([sc], {AGENT_ID: spec}, seed=42)
    agent = spec.build_agent()
    observations = env.reset()
    for _ in range(NUM_STEPS):
        observations, _, _, _ = env.step({AGENT_ID: agent.act(observations[AGENT_ID])})
    obs = observations[AGENT_ID]
    ogm, dagm = obs.occupancy_grid_map, obs.drivable_area_grid_map
    assert ogm.data.shape == dagm.data.shape == (MAP_HEIGHT, MAP_WIDTH, 1)

    def sample_vehicle_pos(pos):
        x, y, _ = project_2d(ogm.metadata, pos)
        assert np.count_nonzero(apply_tolerance(ogm.data, x, y, 2)), ("ogm", pos, x, y)
        x, y, _ = project_2d(dagm.metadata, pos)
        assert np.count_nonzero(apply_tolerance(dagm.data, x, y, 2)), ("drivable area", pos, x, y)

    sample_vehicle_pos(obs.ego_vehicle_state.position)
    assert len(obs.neighborhood_vehicle_states) >= 3          # the test is vacuous without neighbours on the film
    on_film = 0
    for nv in obs.neighborhood_vehicle_states:
        sample_vehicle_pos(nv.position)
        on_film += project_2d(ogm.metadata, nv.position)[2]
    assert on_film >= 3
    n_paths = 0
    for paths in obs.road_waypoints.lanes.values():
        for path in paths:
            for wp in (path[0], path[-1]):
                x, y, inside = project_2d(dagm.metadata, np.array([wp.pos[0], wp.pos[1], 0.0]))
                assert inside and np.count_nonzero(dagm.data[x, y, :]), ("road waypoint off the drivable area", wp.pos, x, y)
            n_paths += 1
    assert n_paths >= 3
    env.close()
