"""The granular terrain co-simulation node (chrono_wisc_b200/terrain.py) against the behaviour of
ChVehicleCosimTerrainNodeGranularOMP ("GOMP": src/chrono_vehicle/cosim/terrain/ChVehicleCosimTerrainNodeGranularOMP.cpp):
settle + checkpoint round trip, the settling observables, rigid and mesh proxies, and the face -> vertex force split."""
import numpy as np
import pytest

from chrono_wisc_b200 import scenes, terrain

pytestmark = pytest.mark.gpu

R = 0.01


def make_node(tmp_path, sampling=terrain.HCP_PACK, depth=0.07):
    node = terrain.ChVehicleCosimTerrainNodeGranularB200(0.4, 0.3, "SMC")
    node.SetOutDir(str(tmp_path))
    node.SetStepSize(1e-4)
    node.SetWallThickness(0.1)
    node.SetGranularMaterial(R, 2500.0)
    node.SetSamplingMethod(sampling, depth, 1.02, False)
    node.SetMaterialSurface(scenes.smc_material(young=5e6, friction=0.5, restitution=0.2))
    node.SetSettlingTime(0.03)
    return node


def test_settle_observables_and_checkpoint_round_trip(tmp_path):
    node = make_node(tmp_path)
    node.Construct()
    n = node.m_num_particles
    assert n > 500
    stats = node.Settle(check_every=100)
    assert stats["Number particles"] == n and stats["Maximum number contacts"] > n
    assert (tmp_path / "settling_stats.info").read_text().startswith("Number particles:")
    # the observables are the reference's formulas (GOMP:700-739)
    p, q, v = node.m_system.GetBodyState(node.m_particles_start, n)
    m = 2500.0 * 4.0 / 3.0 * np.pi * R ** 3
    ke = 0.5 * (m * (v[:, :3] ** 2).sum() + 0.4 * m * R * R * (v[:, 3:] ** 2).sum())
    assert np.isclose(node.CalcTotalKineticEnergy(), ke, rtol=1e-12)
    depth, eta = node.CalculatePackingDensity()
    assert depth == p[:, 2].max() and np.isclose(eta, n * 4 / 3 * np.pi * R ** 3 / (0.4 * 0.3 * depth))
    assert 0.3 < eta < 0.9, eta  # the depth is measured to the highest CENTRE: a shallow HCP bed overshoots 0.74
    assert node.GetInitHeight() == depth + R
    # nobody left the box
    act, _ = node.m_system.GetBodyFlags()
    assert act[1:].all() and not act[0]

    # lossless checkpoint: a second node built the same way and started from it is in the same state
    path = node.WriteCheckpoint("checkpoint_17.dat", precision=17)
    head = open(path).read().split("\n")[:3]
    assert int(head[1]) == n and head[2].split()[0] == str(terrain.TAG_PARTICLES) and len(head[2].split()) == 15
    node2 = make_node(tmp_path)
    node2.SetInputFromCheckpoint("checkpoint_17.dat")
    node2.Construct()
    p2, q2, v2 = node2.m_system.GetBodyState(node2.m_particles_start, n)
    assert np.array_equal(p2, p) and np.array_equal(q2, q)
    assert np.allclose(v2, v, rtol=1e-13, atol=1e-15)  # rot_dt -> w goes through one quaternion product each way
    # and the two then evolve alike
    node.m_system.DoSteps(20)
    node2.m_system.DoSteps(20)
    pa, _, _ = node.m_system.GetBodyState(node.m_particles_start, n)
    pb, _, _ = node2.m_system.GetBodyState(node2.m_particles_start, n)
    assert np.abs(pa - pb).max() < 1e-9
    # the reference's 6-digit format reads back to 6 digits
    node.WriteCheckpoint("checkpoint_6.dat")
    node3 = make_node(tmp_path)
    node3.SetInputFromCheckpoint("checkpoint_6.dat")
    node3.Construct()
    p3, _, _ = node3.m_system.GetBodyState(node3.m_particles_start, n)
    assert np.abs(p3 - pa).max() < 5e-6 * max(1.0, np.abs(pa).max())
    # a checkpoint of another bed is refused
    node4 = make_node(tmp_path, depth=0.05)
    node4.SetInputFromCheckpoint("checkpoint_17.dat")
    with pytest.raises(ValueError, match="inconsistent number of particles"):
        node4.Construct()


def settled_node(tmp_path):
    node = make_node(tmp_path)
    node.SetProxyFixed(True)
    node.Settle(check_every=100)
    return node


def test_rigid_proxy_feels_the_bed(tmp_path):
    node = settled_node(tmp_path)
    top = node.CalcCurrentHeight() + R
    plate = node.AddRigidObject(5.0, boxes=[((0, 0, 0), (0.08, 0.06, 0.01))])
    ball = node.AddRigidObject(1.0, spheres=[((0, 0, 0), 0.03)])
    node.OnInitialize()
    nb0 = node.m_system.num_bodies
    assert nb0 == 1 + node.m_num_particles + 2
    # plate pressed 0.3 r into the top of the bed; the ball overlaps the plate from above (its lowest point is at the
    # top of the bed): proxies are one collision family and do not collide with each other
    node.UpdateRigidProxy(plate, (0.2, 0.0, top + 0.01 - 0.3 * R), (1, 0, 0, 0), (0, 0, 0), (0, 0, 0))
    node.UpdateRigidProxy(ball, (0.2, 0.0, top + 0.03), (1, 0, 0, 0), (0, 0, 0), (0, 0, 1.0))
    node.OnAdvance(2e-4)
    _, f_plate, m_plate = node.GetForceRigidProxy(plate)
    _, f_ball, _ = node.GetForceRigidProxy(ball)
    assert f_plate[2] > 0 and abs(f_plate[0]) < f_plate[2] and abs(f_plate[1]) < f_plate[2]
    c = node.m_system.GetContacts()
    pb, bb = node.m_proxies[plate].first_body, node.m_proxies[ball].first_body
    pair = set(map(tuple, np.sort(c["body_ids"], axis=1)))
    assert (min(pb, bb), max(pb, bb)) not in pair
    # Newton: the plate's force is minus the sum over its contacts of what the particles receive
    f, _ = node.m_system.GetContactForces()
    ids = c["body_ids"]
    on_particles = np.zeros(3)
    for side in (0, 1):
        sel = ids[:, side] == pb
        on_particles += f[1 - side::2][sel].sum(0) if sel.any() else 0.0
    assert np.allclose(on_particles, -f_plate, rtol=1e-9, atol=1e-12)
    # fixed proxies stay where the exchange put them
    p, _, _ = node.m_system.GetBodyState(pb, 1)
    assert np.array_equal(p[0], [0.2, 0.0, top + 0.01 - 0.3 * R])
    # moving the proxy up by a diameter releases the bed
    node.UpdateRigidProxy(plate, (0.2, 0.0, top + 0.01 + 2 * R), (1, 0, 0, 0), (0, 0, 0), (0, 0, 0))
    node.OnAdvance(1e-4)
    assert np.all(node.GetForceRigidProxy(plate)[1] == 0)


def test_mesh_proxy_exchange(tmp_path):
    node = settled_node(tmp_path)
    top = node.CalcCurrentHeight() + R
    # a flat 2 x 1 strip of four triangles over the middle of the bed
    verts = np.array([[0.12, -0.05, 0], [0.20, -0.05, 0], [0.28, -0.05, 0], [0.12, 0.05, 0], [0.20, 0.05, 0], [0.28, 0.05, 0]], float)
    # wound so that the face normals point down, at the bed: triangle_sphere ignores spheres behind a face
    # (ChNarrowphasePRIMS.cpp:398-402), as a tire's outward normals do
    tris = np.array([[0, 4, 1], [0, 3, 4], [1, 5, 2], [1, 4, 5]])
    obj = node.AddMeshObject(verts, tris, 2.0)
    node.OnInitialize()
    px = node.m_proxies[obj]
    assert px.n == 4
    z = top - 0.4 * R
    vpos = verts + [0, 0, z]
    vvel = np.tile([0.1, 0.0, -0.02], (6, 1))
    node.UpdateMeshProxy(obj, vpos, vvel)
    # GOMP:876-905: body at the centroid, identity orientation, mean vertex velocity, no spin
    p, q, v = node.m_system.GetBodyState(px.first_body, 4)
    assert np.allclose(p, vpos[tris].mean(1), rtol=0, atol=1e-16)
    assert np.array_equal(q, np.tile([1.0, 0, 0, 0], (4, 1)))
    assert np.allclose(v[:, :3], [0.1, 0.0, -0.02]) and np.all(v[:, 3:] == 0)
    node.OnAdvance(1e-4)
    vidx, vforce = node.GetForceMeshProxy(obj)
    f, _ = node.m_system.GetBodyContactForces()
    f = f[px.first_body:px.first_body + 4]
    assert np.all(f[:, 2] > 0)  # every face presses on spheres
    assert list(vidx) == [0, 1, 2, 3, 4, 5]
    assert np.allclose(vforce.sum(0), f.sum(0), rtol=1e-12)
    # vertex 0 belongs to faces 0 and 1, vertex 2 to face 2 only: thirds (GOMP:975-1003)
    assert np.allclose(vforce[0], (f[0] + f[1]) / 3, rtol=1e-12) and np.allclose(vforce[2], f[2] / 3, rtol=1e-12)
    # the contacts are with the triangles where the exchange put them, not with the dummy triangles of CreateMeshProxy
    c = node.m_system.GetContacts()
    tri_side = np.isin(c["body_ids"], np.arange(px.first_body, px.first_body + 4))
    pts = np.where(tri_side[:, :1], c["pt_a"], c["pt_b"])[tri_side.any(1)]
    assert pts.shape[0] > 10 and np.abs(pts[:, 2] - z).max() < 1e-9
    assert pts[:, 0].min() > 0.12 - 1e-9 and pts[:, 0].max() < 0.28 + 1e-9
    # lifting half of the strip: only the faces still in the bed report vertices
    vpos2 = vpos.copy()
    vpos2[[2, 5], 2] += 6 * R   # the right-hand faces tilt out of the bed
    vpos2[[1, 4], 2] += 3 * R
    vpos2[[0, 3], 2] += 3 * R
    node.UpdateMeshProxy(obj, vpos2, np.zeros((6, 3)))
    node.OnAdvance(1e-4)
    vidx2, _ = node.GetForceMeshProxy(obj)
    assert vidx2.size == 0
