"""Synthetic scenes for the configurations named in BASELINE.json, as plain numpy arrays.

A scene is a dict with body arrays (``mass, inertia, pos, rot, vel, active, collide``), shape arrays in the
reference's ``shape_container`` layout (``shape_type, shape_body, shape_dims, shape_pos, shape_rot, shape_fam,
shape_mat``), a material table and a ``settings`` dict whose keys are ``cb200_settings`` fields.  The same dict
feeds the CUDA backend (``system.load_scene``), the CPU oracle and the reference build, so all three start from
identical bits.
"""
import numpy as np

from .system import (BOX, CAPSULE, CONVEXHULL, SPHERE, TRIANGLE, FIXED_BIN_SIZE, FIXED_DENSITY, FIXED_RESOLUTION, HERTZ, HYBRID, MODE_SLIDING,
                     SOLVER_APGD)

SMC_MAT_FIELDS = ("young_modulus", "poisson_ratio", "static_friction", "sliding_friction", "rolling_friction",
                  "spinning_friction", "restitution", "constant_adhesion", "adhesionMultDMT", "adhesionSPerko",
                  "kn", "kt", "gn", "gt")
NSC_MAT_FIELDS = ("static_friction", "sliding_friction", "rolling_friction", "spinning_friction", "restitution",
                  "cohesion", "dampingf", "compliance", "complianceT", "complianceRoll", "complianceSpin")


def smc_material(young=2e5, poisson=0.3, friction=0.6, restitution=0.4, adhesion=0.0, rolling=0.0, spinning=0.0,
                 kn=2e5, kt=2e5, gn=40.0, gt=20.0, dmt=0.0, perko=0.0):
    """One row of the SMC material table; defaults of ChContactMaterialSMC
    (src/chrono/physics/ChContactMaterialSMC.cpp:27-37, ChContactMaterial.cpp:32)."""
    return np.array([young, poisson, friction, friction, rolling, spinning, restitution, adhesion, dmt, perko, kn, kt,
                     gn, gt], dtype=np.float32)


def nsc_material(friction=0.6, restitution=0.4, cohesion=0.0, rolling=0.0, spinning=0.0, dampingf=0.0,
                 compliance=0.0, complianceT=0.0, complianceRoll=0.0, complianceSpin=0.0):
    return np.array([friction, friction, rolling, spinning, restitution, cohesion, dampingf, compliance, complianceT,
                     complianceRoll, complianceSpin], dtype=np.float32)


class SceneBuilder:
    def __init__(self):
        self.mass, self.inertia, self.pos, self.rot, self.vel, self.active, self.collide = [], [], [], [], [], [], []
        self.s_type, self.s_body, self.s_dims, self.s_pos, self.s_rot, self.s_fam, self.s_mat = [], [], [], [], [], [], []
        self.nb = 0
        self.verts = []  # vertex pool of triangles (body frame) and convex hulls (shape frame)
        self.nv = 0

    def add_vertices(self, xyz):
        """Append points to the vertex pool; returns the index of the first."""
        xyz = np.asarray(xyz, dtype=np.float64).reshape(-1, 3)
        first = self.nv
        self.verts.append(xyz)
        self.nv += xyz.shape[0]
        return first

    def add_triangle(self, body, v0, v1, v2, **kw):
        """ChCollisionShapeTriangle on `body`: vertices in the body frame (triangle_rigid)."""
        self.add_shapes(TRIANGLE, [body], [[self.add_vertices([v0, v1, v2]), 3, 0, 0]], **kw)

    def add_convex_hull(self, body, points, **kw):
        """ChCollisionShapeConvexHull on `body`: points in the shape frame (convex_rigid)."""
        points = np.asarray(points, dtype=np.float64).reshape(-1, 3)
        self.add_shapes(CONVEXHULL, [body], [[self.add_vertices(points), points.shape[0], 0, 0]], **kw)

    def add_bodies(self, mass, inertia, pos, rot=None, vel=None, active=1, collide=1):
        pos = np.atleast_2d(np.asarray(pos, dtype=np.float64))
        n = pos.shape[0]
        mass = np.broadcast_to(np.asarray(mass, dtype=np.float64), (n,))
        inertia = np.broadcast_to(np.asarray(inertia, dtype=np.float64), (n, 3))
        if rot is None:
            rot = np.tile([1.0, 0, 0, 0], (n, 1))
        if vel is None:
            vel = np.zeros((n, 6))
        first = self.nb
        self.mass.append(mass), self.inertia.append(inertia), self.pos.append(pos)
        self.rot.append(np.asarray(rot, dtype=np.float64).reshape(n, 4))
        self.vel.append(np.asarray(vel, dtype=np.float64).reshape(n, 6))
        self.active.append(np.full(n, active, dtype=np.int8)), self.collide.append(np.full(n, collide, dtype=np.int8))
        self.nb += n
        return first

    def add_shapes(self, type, body, dims, pos=None, rot=None, fam=(1, 0x7FFF), mat=0):
        body = np.atleast_1d(np.asarray(body, dtype=np.uint32))
        n = body.size
        d = np.zeros((n, 4))
        dims = np.atleast_2d(np.asarray(dims, dtype=np.float64))
        dims = np.broadcast_to(dims, (n, dims.shape[1]))
        d[:, :dims.shape[1]] = dims
        self.s_type.append(np.broadcast_to(np.asarray(type, dtype=np.int32), (n,)))
        self.s_body.append(body), self.s_dims.append(d)
        self.s_pos.append(np.zeros((n, 3)) if pos is None else np.asarray(pos, dtype=np.float64).reshape(n, 3))
        self.s_rot.append(np.tile([1.0, 0, 0, 0], (n, 1)) if rot is None else np.asarray(rot, dtype=np.float64).reshape(n, 4))
        self.s_fam.append(np.broadcast_to(np.asarray(fam, dtype=np.int16), (n, 2)))
        self.s_mat.append(np.broadcast_to(np.asarray(mat, dtype=np.int32), (n,)))

    def add_box_container(self, body, center, size, thickness, faces=(2, 2, -1), mat=0):
        """utils::AddBoxContainer (src/chrono/utils/ChUtilsCreators.cpp:559-615): walls are boxes attached to `body`;
        order Z-, Z+, X-, X+, Y-, Y+ as in the reference."""
        sx, sy, sz = size
        hx, hy, hz, ht = sx / 2, sy / 2, sz / 2, thickness / 2
        c = np.asarray(center, dtype=np.float64)
        walls = []
        if faces[2] in (-1, 2):
            walls.append(((0, 0, -hz - ht), (sx / 2, sy / 2, ht)))
        if faces[2] in (1, 2):
            walls.append(((0, 0, hz + ht), (sx / 2, sy / 2, ht)))
        if faces[0] in (-1, 2):
            walls.append(((-hx - ht, 0, 0), (ht, sy / 2, sz / 2)))
        if faces[0] in (1, 2):
            walls.append(((hx + ht, 0, 0), (ht, sy / 2, sz / 2)))
        if faces[1] in (-1, 2):
            walls.append(((0, -hy - ht, 0), (sx / 2, ht, sz / 2)))
        if faces[1] in (1, 2):
            walls.append(((0, hy + ht, 0), (sx / 2, ht, sz / 2)))
        for p, h in walls:
            self.add_shapes(BOX, [body], np.array([h]), pos=[c + np.asarray(p)], mat=mat)

    def finish(self, materials, settings):
        cat = np.concatenate
        return dict(
            mass=cat(self.mass), inertia=cat(self.inertia), pos=cat(self.pos), rot=cat(self.rot), vel=cat(self.vel),
            active=cat(self.active), collide=cat(self.collide), shape_type=cat(self.s_type).astype(np.int32),
            shape_body=cat(self.s_body).astype(np.uint32), shape_dims=cat(self.s_dims), shape_pos=cat(self.s_pos),
            shape_rot=cat(self.s_rot), shape_fam=cat(self.s_fam).astype(np.int16),
            shape_mat=cat(self.s_mat).astype(np.int32), materials=np.atleast_2d(np.asarray(materials, np.float32)),
            shape_verts=cat(self.verts) if self.verts else np.zeros((0, 3)), settings=dict(settings))


def lattice_positions(nx, ny, nz, spacing, jitter, z0, seed):
    """nx*ny*nz sphere centres on a jittered lattice centred on the z axis; x fastest, then y, then z."""
    rng = np.random.Generator(np.random.MT19937(seed))
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    p = np.stack([ix.transpose(2, 1, 0).ravel(), iy.transpose(2, 1, 0).ravel(), iz.transpose(2, 1, 0).ravel()], 1)
    p = p.astype(np.float64)
    p[:, 0] = (p[:, 0] - (nx - 1) / 2) * spacing
    p[:, 1] = (p[:, 1] - (ny - 1) / 2) * spacing
    p[:, 2] = z0 + p[:, 2] * spacing
    p += rng.uniform(-jitter, jitter, size=p.shape)
    return p


def settling_smc(nx=32, ny=32, nz=8, radius=0.02, seed=12345, grid="density", bins=(10, 10, 1), bin_edge=2.0, step=1e-3,
                 young=2e6, mu=0.4, cr=0.4, rho=2000.0, spacing_factor=2.02, jitter_factor=0.005, wall_thickness=0.2):
    """SMC Hertz settling of monodisperse spheres in an open box (configs A/B/C of BASELINE.json; scene after
    src/tests/benchmark_tests/multicore/btest_MCORE_settling.cpp:59-131 with a jittered lattice instead of the
    Poisson-disk sampler).  1M spheres = settling_smc(250, 250, 16)."""
    b = SceneBuilder()
    spacing = spacing_factor * radius
    sx = nx * spacing + 2 * radius * 0.5
    sy = ny * spacing + 2 * radius * 0.5
    sz = max(1.0, (nz + 2) * spacing)
    bin_body = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
    b.add_box_container(bin_body, (0, 0, sz / 2), (sx, sy, sz), wall_thickness, faces=(2, 2, -1))
    m = rho * 4.0 / 3.0 * np.pi * radius ** 3
    inertia = 0.4 * m * radius * radius
    pos = lattice_positions(nx, ny, nz, spacing, jitter_factor * radius, 1.01 * radius, seed)
    first = b.add_bodies(m, (inertia, inertia, inertia), pos)
    n = pos.shape[0]
    b.add_shapes(SPHERE, np.arange(first, first + n), np.full((n, 1), radius))
    settings = dict(system_type=1, step_size=step, gravity=(0.0, 0.0, -9.81), contact_force_model=HERTZ,
                    adhesion_force_model=0, tangential_displ_mode=1, use_material_properties=1,
                    narrowphase_algorithm=HYBRID, collision_envelope=0.0)
    if grid == "density":
        settings.update(broadphase_grid=FIXED_DENSITY, grid_density=5.0)
    elif grid == "bin_size":
        # FIXED_BIN_SIZE with a cubic bin of `bin_edge` sphere diameters ("or bin edge 2.5 diameters")
        settings.update(broadphase_grid=FIXED_BIN_SIZE, bin_size=(bin_edge * 2 * radius,) * 3)
    else:
        settings.update(broadphase_grid=FIXED_RESOLUTION, bins_per_axis=tuple(bins))
    return b.finish([smc_material(young=young, friction=mu, restitution=cr)], settings)


def mixer_smc(nx=24, ny=24, nz=6, radius=0.02, seed=5, step=5e-4, blade=(0.3, 0.02, 0.06), blade_mass=5.0):
    """demo_MCORE_mixerSMC-style scene (config E flavour): a sphere bed in a box with a blade (a box body) standing in
    it, to be driven about the vertical axis by a speed motor (links.motor_rotation_speed_z).  Returns the scene, the
    index of the blade body and its anchor point."""
    b = SceneBuilder()
    spacing = 2.02 * radius
    sx, sy = nx * spacing + radius, ny * spacing + radius
    sz = max(0.5, (nz + 4) * spacing)
    bin_body = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
    b.add_box_container(bin_body, (0, 0, sz / 2), (sx, sy, sz), 0.2, faces=(2, 2, -1))
    m = 2000.0 * 4.0 / 3.0 * np.pi * radius ** 3
    pos = lattice_positions(nx, ny, nz, spacing, 0.005 * radius, 1.01 * radius, seed)
    hx, hy, hz = blade
    anchor = np.array([0.0, 0.0, 1.01 * radius + hz + 0.5 * radius])
    # leave room for the blade: drop the spheres its swept disc would start inside of
    keep = ~((np.hypot(pos[:, 0], pos[:, 1]) < hx + 1.5 * radius) & (pos[:, 2] < anchor[2] + hz + 1.5 * radius))
    pos = pos[keep]
    first = b.add_bodies(m, (0.4 * m * radius * radius,) * 3, pos)
    n = pos.shape[0]
    b.add_shapes(SPHERE, np.arange(first, first + n), np.full((n, 1), radius))
    ib = blade_mass / 3.0 * np.array([hy * hy + hz * hz, hx * hx + hz * hz, hx * hx + hy * hy])
    blade_body = b.add_bodies(blade_mass, tuple(ib), [anchor])
    b.add_shapes(BOX, [blade_body], np.array([[hx, hy, hz]]))
    settings = dict(system_type=1, step_size=step, gravity=(0.0, 0.0, -9.81), contact_force_model=HERTZ,
                    adhesion_force_model=0, tangential_displ_mode=1, use_material_properties=1,
                    narrowphase_algorithm=HYBRID, collision_envelope=0.0, broadphase_grid=FIXED_BIN_SIZE,
                    bin_size=(2.5 * 2 * radius,) * 3)
    return b.finish([smc_material(young=2e6, friction=0.4, restitution=0.4)], settings), blade_body, anchor


def mixed_smc(n_side=6, seed=7, radius=0.1, step=1e-3, shapes=("sphere", "box", "capsule")):
    """Small mixed-shape SMC scene (config E flavour, after src/demos/multicore/demo_MCORE_mixerSMC.cpp:51-169):
    a lattice cycling through the requested shape kinds inside an open box."""
    b = SceneBuilder()
    spacing = 2.4 * radius
    sx = sy = n_side * spacing + radius
    sz = (n_side + 2) * spacing
    bin_body = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
    b.add_box_container(bin_body, (0, 0, sz / 2), (sx, sy, sz), 0.2, faces=(2, 2, -1))
    pos = lattice_positions(n_side, n_side, n_side, spacing, 0.02 * radius, 1.2 * radius, seed)
    rng = np.random.Generator(np.random.MT19937(seed + 1))
    n = pos.shape[0]
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    kinds = [shapes[i % len(shapes)] for i in range(n)]
    rho = 1000.0
    for i, k in enumerate(kinds):
        if k == "sphere":
            m = rho * 4 / 3 * np.pi * radius ** 3
            I = (0.4 * m * radius ** 2,) * 3
            bid = b.add_bodies(m, I, [pos[i]], rot=[q[i]])
            b.add_shapes(SPHERE, [bid], [[radius]])
        elif k == "box":
            h = np.array([radius, 0.8 * radius, 0.6 * radius])
            m = rho * 8 * h.prod()
            I = m / 3 * np.array([h[1] ** 2 + h[2] ** 2, h[0] ** 2 + h[2] ** 2, h[0] ** 2 + h[1] ** 2])
            bid = b.add_bodies(m, I, [pos[i]], rot=[q[i]])
            b.add_shapes(BOX, [bid], [h])
        else:
            r, hl = 0.6 * radius, 0.6 * radius  # stored (radius, radius): SURVEY Appendix A.1
            m = rho * (np.pi * r * r * 2 * hl + 4 / 3 * np.pi * r ** 3)
            I = (0.5 * m * r * r,) * 3
            bid = b.add_bodies(m, I, [pos[i]], rot=[q[i]])
            b.add_shapes(CAPSULE, [bid], [[r, hl]])
    settings = dict(system_type=1, step_size=step, gravity=(0.0, 0.0, -9.81), contact_force_model=HERTZ,
                    narrowphase_algorithm=HYBRID, broadphase_grid=FIXED_DENSITY, grid_density=5.0)
    return b.finish([smc_material(young=2e5, friction=0.4, restitution=0.1)], settings)


def mixed_bed(nx=180, ny=180, nz=16, radius=0.1, seed=7, step=1e-3, algorithm=HYBRID, bin_edge=3.0):
    """Large mixed-shape SMC bed (BASELINE.json configs[4] flavour, shape families of
    src/demos/multicore/demo_MCORE_mixerSMC.cpp:118-168): one third spheres (r), boxes (r, 0.8 r, 0.6 r) and capsules
    (0.6 r, stored half-length as Appendix A.1 of the survey), E = 2e5, mu = 0.4, cr = 0.1, settling in an open box;
    FIXED_BIN_SIZE with a cubic bin of `bin_edge` radii.  518,400 bodies = mixed_bed(180, 180, 16).  The mixer blade of
    the demo is a motor-driven body (bilateral constraint) and is not part of this scene."""
    b = SceneBuilder()
    rho = 1000.0
    spacing = 2.4 * radius
    sx, sy, sz = nx * spacing + radius, ny * spacing + radius, (nz + 2) * spacing
    bin_body = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
    b.add_box_container(bin_body, (0, 0, sz / 2), (sx, sy, sz), 0.2, faces=(2, 2, -1))
    pos = lattice_positions(nx, ny, nz, spacing, 0.02 * radius, 1.2 * radius, seed)
    rng = np.random.Generator(np.random.MT19937(seed + 1))
    n = pos.shape[0]
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    kind = np.arange(n) % 3
    for k in range(3):
        sel = np.flatnonzero(kind == k)
        if k == 0:
            m = rho * 4 / 3 * np.pi * radius ** 3
            I = (0.4 * m * radius ** 2,) * 3
            first = b.add_bodies(m, I, pos[sel], rot=q[sel])
            b.add_shapes(SPHERE, first + np.arange(sel.size), [[radius]])
        elif k == 1:
            h = np.array([radius, 0.8 * radius, 0.6 * radius])
            m = rho * 8 * h.prod()
            I = m / 3 * np.array([h[1] ** 2 + h[2] ** 2, h[0] ** 2 + h[2] ** 2, h[0] ** 2 + h[1] ** 2])
            first = b.add_bodies(m, I, pos[sel], rot=q[sel])
            b.add_shapes(BOX, first + np.arange(sel.size), [h])
        else:
            r, hl = 0.6 * radius, 0.6 * radius
            m = rho * (np.pi * r * r * 2 * hl + 4 / 3 * np.pi * r ** 3)
            I = (0.5 * m * r * r,) * 3
            first = b.add_bodies(m, I, pos[sel], rot=q[sel])
            b.add_shapes(CAPSULE, first + np.arange(sel.size), [[r, hl]])
    settings = dict(system_type=1, step_size=step, gravity=(0.0, 0.0, -9.81), contact_force_model=HERTZ,
                    narrowphase_algorithm=algorithm, broadphase_grid=FIXED_BIN_SIZE, bin_size=(bin_edge * radius,) * 3)
    return b.finish([smc_material(young=2e5, friction=0.4, restitution=0.1)], settings)


def poisson_disk_2d(half_x, half_y, min_dist, rng, k=30):
    """Poisson-disk sample of the rectangle [-half_x, half_x] x [-half_y, half_y]: no two points closer than min_dist,
    dart throwing around active points with a background grid (Bridson) -- the scheme of utils::ChPDSampler
    (src/chrono/utils/ChUtilsSamplers.h:264-469: grid of cell size min_dist / sqrt(2), up to 30 candidates in the annulus
    [d, 2d] around an active point).  Not the reference's random stream: both arms of a comparison get the same sample."""
    cell = min_dist / np.sqrt(2.0)
    nx, ny = int(np.ceil(2 * half_x / cell)) + 1, int(np.ceil(2 * half_y / cell)) + 1
    grid = -np.ones((nx, ny), dtype=np.int64)
    pts = []

    def put(p):
        grid[int((p[0] + half_x) / cell), int((p[1] + half_y) / cell)] = len(pts)
        pts.append(p)

    put(np.array([rng.uniform(-half_x, half_x), rng.uniform(-half_y, half_y)]))
    active = [0]
    d2 = min_dist * min_dist
    while active:
        ai = int(rng.integers(len(active)))
        base = pts[active[ai]]
        rad = min_dist * np.sqrt(rng.uniform(1.0, 4.0, k))
        ang = rng.uniform(0.0, 2 * np.pi, k)
        cand = base + np.stack([rad * np.cos(ang), rad * np.sin(ang)], 1)
        placed = False
        for c in cand:
            if abs(c[0]) > half_x or abs(c[1]) > half_y:
                continue
            gx, gy = int((c[0] + half_x) / cell), int((c[1] + half_y) / cell)
            nb = grid[max(gx - 2, 0):gx + 3, max(gy - 2, 0):gy + 3].ravel()
            nb = nb[nb >= 0]
            if nb.size and (((np.asarray([pts[j] for j in nb]) - c) ** 2).sum(1) < d2).any():
                continue
            put(c)
            active.append(len(pts) - 1)
            placed = True
            break
        if not placed:
            active[ai] = active[-1]
            active.pop()
    return np.asarray(pts)


def settling_pd(half=(2.0, 2.0), layers=8, radius=0.02, seed=0, step=1e-3, bins=(10, 10, 1)):
    """BASELINE.json configs[0], the reference's own settling benchmark scene
    (src/tests/benchmark_tests/multicore/btest_MCORE_settling.cpp:59-131): a fixed bin of five boxes
    (AddBoxContainer, frame (0, 0, 0.5), size (2 hx, 2 hy, 1), thickness 0.2, faces (2, 2, -1)); spheres r = 0.02,
    rho = 2000, E = 2e6, mu = 0.4, cr = 0.4, no adhesion, in `layers` layers at z = 2r', 4r', ... (r' = 1.01 r), each a
    Poisson-disk sample with minimum distance 2r' over +-(h - r'); Hertz, OneStep, material properties, HYBRID
    narrowphase, FIXED_RESOLUTION bins_per_axis = 10 x 10 x 1.  half = (2, 2), 8 layers is the reference's size (~55 k
    spheres); smaller footprints give the "~10k" variants."""
    b = SceneBuilder()
    bin_body = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
    b.add_box_container(bin_body, (0, 0, 0.5), (2 * half[0], 2 * half[1], 1.0), 0.2, faces=(2, 2, -1))
    rp = 1.01 * radius
    rng = np.random.default_rng(seed)
    pos = []
    for il in range(layers):
        xy = poisson_disk_2d(half[0] - rp, half[1] - rp, 2 * rp, rng)
        pos.append(np.concatenate([xy, np.full((xy.shape[0], 1), 2 * rp * (il + 1))], 1))
    pos = np.concatenate(pos)
    m = 2000.0 * 4.0 / 3.0 * np.pi * radius ** 3
    first = b.add_bodies(m, (0.4 * m * radius * radius,) * 3, pos)
    n = pos.shape[0]
    b.add_shapes(SPHERE, np.arange(first, first + n), np.full((n, 1), radius))
    settings = dict(system_type=1, step_size=step, gravity=(0.0, 0.0, -9.81), contact_force_model=HERTZ,
                    adhesion_force_model=0, tangential_displ_mode=1, use_material_properties=1,
                    narrowphase_algorithm=HYBRID, collision_envelope=0.0, broadphase_grid=FIXED_RESOLUTION,
                    bins_per_axis=tuple(bins), tolerance=1e-3)
    return b.finish([smc_material(young=2e6, friction=0.4, restitution=0.4)], settings)


def contact_force_test(method="nsc"):
    """The scene of the reference's contact-force acceptance test (src/tests/unit_tests/multicore/
    utest_MCORE_contact_forces.cpp:42-152): 8 balls (r = 0.5, m = 5) on a diagonal at 1.1 r above the floor of a
    20 x 20 x 2r box container, h = 1e-3; NSC: SLIDING, APGD, 100 iterations, tolerance 1e-5, mu 0.4, cr 0; SMC: Hooke
    with user coefficients kn = 2e4, gn = 5e2, kt = gt = 0, OneStep.  Once at rest the contact force on the container
    equals the weight of the balls (rel 1e-3 in the reference, checked from t = 0.5 s to 3 s)."""
    b = SceneBuilder()
    r, m, n = 0.5, 5.0, 8
    ground = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
    b.add_box_container(ground, (0, 0, r), (20, 20, 2 * r), 0.1, faces=(2, 2, -1))
    pos = np.array([[i * 2 * r, i * 2 * r, 1.1 * r] for i in range(n)]) - np.array([7 * r, 7 * r, 0.0])
    first = b.add_bodies(m, (0.4 * m * r * r,) * 3, pos)
    b.add_shapes(SPHERE, np.arange(first, first + n), np.full((n, 1), r))
    common = dict(step_size=1e-3, gravity=(0.0, 0.0, -9.81), narrowphase_algorithm=HYBRID, broadphase_grid=FIXED_RESOLUTION,
                  bins_per_axis=(10, 10, 10), tolerance=1e-5)
    if method == "nsc":
        settings = dict(common, system_type=0, collision_envelope=0.0, solver_type=SOLVER_APGD, solver_mode=MODE_SLIDING,
                        max_iteration_normal=0, max_iteration_sliding=100, max_iteration_spinning=0)
        return b.finish([nsc_material(friction=0.4, restitution=0.0)], settings)
    settings = dict(common, system_type=1, contact_force_model=0, tangential_displ_mode=1, use_material_properties=0)
    return b.finish([smc_material(young=2e4, friction=0.4, restitution=0.0, kn=2e4, gn=5e2, kt=0.0, gt=0.0)], settings)


def balls_nsc(nx=16, ny=16, nz=4, radius=0.15, seed=3, step=1e-2, envelope=0.01, iterations=100, mu=0.4):
    """NSC ball bed (config D, after src/demos/multicore/demo_MCORE_ballsNSC.cpp:121-149): unit-mass spheres with
    I = 0.4 m r^2, friction 0.4, SLIDING mode, alpha = 0, recovery speed 1e4, APGD with a fixed iteration count.
    262,144 spheres = balls_nsc(128, 128, 16)."""
    b = SceneBuilder()
    spacing = 2.05 * radius
    sx = nx * spacing + radius
    sy = ny * spacing + radius
    sz = (nz + 4) * spacing
    bin_body = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
    b.add_box_container(bin_body, (0, 0, sz / 2), (sx, sy, sz), 0.2, faces=(2, 2, -1))
    m = 1.0
    inertia = 0.4 * m * radius * radius
    pos = lattice_positions(nx, ny, nz, spacing, 0.01 * radius, 1.05 * radius, seed)
    first = b.add_bodies(m, (inertia,) * 3, pos)
    n = pos.shape[0]
    b.add_shapes(SPHERE, np.arange(first, first + n), np.full((n, 1), radius))
    settings = dict(system_type=0, step_size=step, gravity=(0.0, 0.0, -9.81), collision_envelope=envelope,
                    narrowphase_algorithm=HYBRID, broadphase_grid=FIXED_DENSITY, grid_density=5.0,
                    solver_type=SOLVER_APGD, solver_mode=MODE_SLIDING, max_iteration_normal=0,
                    max_iteration_sliding=iterations, max_iteration_spinning=0, alpha=0.0,
                    contact_recovery_speed=1e4, tolerance=0.0)
    return b.finish([nsc_material(friction=mu)], settings)


def settling_smc_slab(rank, world, nx=250, ny=250, nz=16, radius=0.02, seed=12345, bin_edge=2.5, step=1e-3,
                      young=2e6, mu=0.4, cr=0.4, rho=2000.0, spacing_factor=2.0, jitter_factor=0.005,
                      wall_thickness=0.2, ghost_width=None):
    """Rank ``rank``'s share of the (world * nx) x ny x nz settling bed of BASELINE.json configs[2] (8M spheres =
    8 ranks of 250 x 250 x 16), generated locally: slabs are cut along x between lattice planes; the container is one
    static body whose floor and y walls are boxes covering this rank's interval plus a margin, the x walls exist on
    the first and last rank only.  Returns (scene, gids, cuts, ghost_width); gids follow the global lattice order
    (x fastest, then y, then z; body 0 is the container)."""
    b = SceneBuilder()
    spacing = spacing_factor * radius
    gnx = world * nx
    sx = gnx * spacing + 2 * radius * 0.5
    sy = ny * spacing + 2 * radius * 0.5
    sz = max(1.0, (nz + 2) * spacing)
    if ghost_width is None:
        ghost_width = 2 * radius * 1.25  # >= 2 r_max + 2 envelope (SMC: envelope 0), with room for one step of motion
    x_of = lambda i: (i - (gnx - 1) / 2) * spacing
    cuts = [x_of(g * nx) - spacing / 2 for g in range(world + 1)]
    cuts[0], cuts[-1] = -sx, sx
    lo = -sx / 2 if rank == 0 else cuts[rank] - 4 * ghost_width
    hi = sx / 2 if rank == world - 1 else cuts[rank + 1] + 4 * ghost_width
    cx, lx = (lo + hi) / 2, hi - lo
    ht = wall_thickness / 2
    bin_body = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
    walls = [((cx, 0, -ht), (lx / 2, sy / 2, ht)),
             ((cx, -sy / 2 - ht, sz / 2), (lx / 2, ht, sz / 2)), ((cx, sy / 2 + ht, sz / 2), (lx / 2, ht, sz / 2))]
    if rank == 0:
        walls.append(((-sx / 2 - ht, 0, sz / 2), (ht, sy / 2, sz / 2)))
    if rank == world - 1:
        walls.append(((sx / 2 + ht, 0, sz / 2), (ht, sy / 2, sz / 2)))
    for p, h in walls:
        b.add_shapes(BOX, [bin_body], np.array([h]), pos=[np.asarray(p, dtype=np.float64)])
    m = rho * 4.0 / 3.0 * np.pi * radius ** 3
    inertia = 0.4 * m * radius * radius
    rng = np.random.Generator(np.random.MT19937(seed + rank))
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    ix, iy, iz = (a.transpose(2, 1, 0).ravel() for a in (ix, iy, iz))
    gi = ix + rank * nx
    pos = np.stack([x_of(gi), (iy - (ny - 1) / 2) * spacing, 1.01 * radius + iz * spacing], 1).astype(np.float64)
    pos += rng.uniform(-jitter_factor * radius, jitter_factor * radius, size=pos.shape)
    first = b.add_bodies(m, (inertia, inertia, inertia), pos)
    n = pos.shape[0]
    b.add_shapes(SPHERE, np.arange(first, first + n), np.full((n, 1), radius))
    gids = (iz.astype(np.int64) * ny + iy) * gnx + gi + 1
    settings = dict(system_type=1, step_size=step, gravity=(0.0, 0.0, -9.81), contact_force_model=HERTZ,
                    adhesion_force_model=0, tangential_displ_mode=1, use_material_properties=1,
                    narrowphase_algorithm=HYBRID, collision_envelope=0.0, broadphase_grid=FIXED_BIN_SIZE,
                    bin_size=(bin_edge * 2 * radius,) * 3)
    scene = b.finish([smc_material(young=young, friction=mu, restitution=cr)], settings)
    return scene, gids, cuts, ghost_width
