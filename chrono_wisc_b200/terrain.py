"""Granular terrain node of the vehicle co-simulation framework on top of the B200 hot path.

Mirror of ``ChVehicleCosimTerrainNodeGranularOMP``
(src/chrono_vehicle/cosim/terrain/ChVehicleCosimTerrainNodeGranularOMP.cpp, "GOMP" below): the main real-world caller
of Chrono::Multicore -- a box of granular material, settled once and checkpointed, that exchanges states and forces with
a vehicle's tires through PROXY bodies (one triangle body per mesh face for deformable tires, one body for rigid ones).
Same method names, same order of operations, same file formats.  What is left out is the MPI plumbing of the base
classes (ChVehicleCosimBaseNode / ...TerrainNode: the exchange with the MBS and tire nodes): states come in and forces
go out through plain method arguments, so the node can be driven by any transport.  No rendering (no graphics stack:
unavailable on this box).

Host code only; every step runs through the C ABI (``ChSystemMulticoreSMC`` / ``NSC`` in system.py).
"""
import os

import numpy as np

from . import scenes
from .scenes import BOX, SPHERE, TRIANGLE, SceneBuilder
from .system import ChSystemMulticoreNSC, ChSystemMulticoreSMC, load_scene

POISSON_DISK, HCP_PACK, REGULAR_GRID = 0, 1, 2  # utils::SamplingType

# body tags (GOMP:53-55): container -1, obstacles 100000, particles from 200000
TAG_OBSTACLES, TAG_PARTICLES = 100000, 200000


def grid_points(center, hdims, sep):
    """utils::ChGridSampler::Sample on a box (src/chrono/utils/ChUtilsSamplers.h:501-524)."""
    center, hdims = np.asarray(center, float), np.asarray(hdims, float)
    d = 2 * hdims
    n = np.where(d > 0, np.round(d / sep), 0).astype(int)
    s = np.where(n > 0, d / np.maximum(n, 1), 0.0)
    bl = center - hdims
    ix, iy, iz = np.meshgrid(np.arange(n[0] + 1), np.arange(n[1] + 1), np.arange(n[2] + 1), indexing="ij")
    return bl + np.stack([ix.ravel() * s[0], iy.ravel() * s[1], iz.ravel() * s[2]], 1)


def hcp_points(center, hdims, sep):
    """utils::ChHCPSampler::Sample on a box (ChUtilsSamplers.h:540-569); points outside the box are rejected."""
    center, hdims = np.asarray(center, float), np.asarray(hdims, float)
    bl = center - hdims
    dx, dy, dz = sep, sep * np.sqrt(3.0) / 2, sep * np.sqrt(2.0 / 3.0)
    nx, ny, nz = int(2 * hdims[0] / dx) + 1, int(2 * hdims[1] / dy) + 1, int(2 * hdims[2] / dz) + 1
    out = []
    for k in range(nz):
        oy = 0.0 if k % 2 == 0 else dy / 3
        for j in range(ny):
            ox = 0.0 if (j + k) % 2 == 0 else dx / 2
            p = bl + np.stack([ox + np.arange(nx) * dx, np.full(nx, oy + j * dy), np.full(nx, k * dz)], 1)
            out.append(p)
    p = np.concatenate(out)
    ok = np.all(np.abs(p - center) <= hdims + 1e-12, axis=1)
    return p[ok]


class ProxyBodySet:
    """GOMP.h:149-154: the bodies that stand for one co-simulated object, and where their triangles live."""

    def __init__(self, first_body, n, first_vertex=None, idx_verts=None):
        self.first_body, self.n = first_body, n
        self.first_vertex, self.idx_verts = first_vertex, idx_verts


class ChVehicleCosimTerrainNodeGranularB200:
    """Terrain patch of length (X) x width (Y); the patch spans x in [0, length], y in [-width/2, width/2] as in GOMP."""

    def __init__(self, length, width, method="SMC", device=0):
        self.m_dimX, self.m_dimY = float(length), float(width)
        self.m_method = method.upper()
        self.m_system = (ChSystemMulticoreSMC if self.m_method == "SMC" else ChSystemMulticoreNSC)(device)
        # defaults of the reference's constructor (GOMP:57-121)
        self.m_thick = 0.2
        self.m_radius_g, self.m_rho_g = 0.01, 2000.0
        self.m_sampling_type, self.m_init_depth, self.m_separation_factor, self.m_in_layers = POISSON_DISK, 0.2, 1.001, False
        self.m_material = None
        self.m_use_checkpoint, self.m_checkpoint_filename = False, ""
        self.m_fixed_settling_duration, self.m_time_settling, self.m_KE_settling = True, 0.4, 1e-3
        self.m_fixed_proxies, self.m_radius_p = False, 5e-3
        self.m_step_size = 1e-4
        self.m_node_out_dir = "."
        self.m_constructed = False
        self.m_num_particles, self.m_particles_start = 0, 0
        self.m_init_height = 0.0
        self.m_proxies, self.m_objects = {}, []
        self.m_seed = 0
        st = self.m_system.GetSettings()
        st.gravity[:] = [0.0, 0.0, -9.81]
        st.narrowphase_algorithm = scenes.HYBRID
        st.tolerance = 0.1 if self.m_method == "NSC" else st.tolerance
        if self.m_method == "SMC":
            st.contact_force_model, st.tangential_displ_mode, st.use_material_properties = scenes.HERTZ, 1, 1
        else:
            st.collision_envelope = 0.001
            st.max_iteration_normal, st.max_iteration_sliding, st.max_iteration_spinning = 0, 100, 0
            st.contact_recovery_speed, st.alpha = 1e4, 0.0

    # ---- setters, named as in GOMP:261-325 ----------------------------------------------------------------------
    def SetOutDir(self, path):
        self.m_node_out_dir = path
        os.makedirs(path, exist_ok=True)

    def SetStepSize(self, h):
        self.m_step_size = float(h)

    def SetWallThickness(self, thickness):
        self.m_thick = float(thickness)

    def SetGranularMaterial(self, radius, density):
        self.m_radius_g, self.m_rho_g = float(radius), float(density)
        self.m_system.GetSettings().collision_envelope = 0.0 if self.m_method == "SMC" else 0.1 * radius

    def UseMaterialProperties(self, flag):
        self.m_system.GetSettings().use_material_properties = int(bool(flag))

    def SetContactForceModel(self, model):
        self.m_system.GetSettings().contact_force_model = int(model)

    def SetTangentialDisplacementModel(self, model):
        self.m_system.GetSettings().tangential_displ_mode = int(model)

    def SetSamplingMethod(self, type, init_height, sep_factor=1.001, in_layers=False):
        self.m_sampling_type, self.m_init_depth = type, float(init_height)
        self.m_separation_factor, self.m_in_layers = float(sep_factor), bool(in_layers)

    def SetMaterialSurface(self, mat):
        """One material row: scenes.smc_material(...) / scenes.nsc_material(...)."""
        self.m_material = np.asarray(mat, dtype=np.float32)

    def SetProxyFixed(self, fixed):
        self.m_fixed_proxies = bool(fixed)

    def SetProxyContactRadius(self, radius):
        self.m_radius_p = float(radius)

    def SetInputFromCheckpoint(self, filename):
        self.m_use_checkpoint, self.m_checkpoint_filename = True, filename

    def SetSettlingTime(self, time):
        self.m_time_settling = max(float(time), 1e-5)
        self.m_fixed_settling_duration = True

    def SetSettlingKineticEneryThreshold(self, threshold):  # [sic] the reference's spelling
        self.m_KE_settling = float(threshold)
        self.m_fixed_settling_duration = False

    def AddMeshObject(self, vertices, triangles, mass):
        """A co-simulated object with a triangle-mesh contact geometry (what the MBS / tire node sends at
        initialisation: ChVehicleCosimTerrainNode::Initialize).  Returns the object index."""
        self.m_objects.append(dict(kind="mesh", verts=np.asarray(vertices, float).reshape(-1, 3),
                                   tris=np.asarray(triangles, np.int64).reshape(-1, 3), mass=float(mass)))
        return len(self.m_objects) - 1

    def AddRigidObject(self, mass, spheres=None, boxes=None):
        """A rigid co-simulated object: contact spheres [(centre, radius)] and boxes [(centre, half dims)] in the
        object frame."""
        self.m_objects.append(dict(kind="rigid", mass=float(mass), spheres=spheres or [], boxes=boxes or []))
        return len(self.m_objects) - 1

    # ---- construction (GOMP:329-582) ------------------------------------------------------------------------------
    def _sample(self, center, hdims, sep, rng):
        if self.m_sampling_type == REGULAR_GRID:
            return grid_points(center, hdims, sep)
        if self.m_sampling_type == HCP_PACK:
            return hcp_points(center, hdims, sep)
        # POISSON_DISK: a 2-D sample per layer, the layers `sep` apart (ChPDSampler on a box of zero height is what
        # in_layers uses; for a box with height the reference samples in 3-D -- here: stacked 2-D samples)
        nl = max(int(np.floor(2 * hdims[2] / sep)) + 1, 1)
        out = []
        for il in range(nl):
            xy = scenes.poisson_disk_2d(hdims[0], hdims[1], sep, rng)
            out.append(np.concatenate([xy + np.asarray(center[:2]), np.full((xy.shape[0], 1), center[2] - hdims[2] + il * sep)], 1))
        return np.concatenate(out)

    def Construct(self):
        if self.m_constructed:
            return
        if self.m_material is None:
            self.m_material = scenes.smc_material() if self.m_method == "SMC" else scenes.nsc_material()
        r = self.m_separation_factor * self.m_radius_g
        delta = 2.0 * r
        st = self.m_system.GetSettings()
        st.step_size = self.m_step_size
        # broad-phase bins (:350-356)
        factor = 2
        st.broadphase_grid = scenes.FIXED_RESOLUTION
        st.bins_per_axis[:] = [max(int(np.ceil(self.m_dimX / (2 * self.m_radius_g))) // factor, 1),
                               max(int(np.ceil(self.m_dimY / (2 * self.m_radius_g))) // factor, 1), 1]
        b = SceneBuilder()
        # container: bottom, front, rear, left, right boxes on one fixed body (:362-387); AddBoxGeometry takes full sizes
        cont = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
        hX, hY, hZ, ht = self.m_dimX / 2, self.m_dimY / 2, 0.5 * self.m_init_depth, self.m_thick / 2
        D, T = self.m_init_depth, self.m_thick
        walls = [((self.m_dimX, self.m_dimY, T), (hX, 0, -ht)),
                 ((T, self.m_dimY, D + T), (self.m_dimX + ht, 0, hZ - ht)),
                 ((T, self.m_dimY, D + T), (-ht, 0, hZ - ht)),
                 ((self.m_dimX, T, D + T), (hX, hY + ht, hZ - ht)),
                 ((self.m_dimX, T, D + T), (hX, -hY - ht, hZ - ht))]
        for size, pos in walls:
            b.add_shapes(BOX, [cont], [np.asarray(size) / 2], pos=[pos])
        # bodies that leave this box are deactivated (:389-393)
        self.m_active_box = (np.array([-ht, -hY - ht, -ht]), np.array([self.m_dimX + ht, hY + ht, 2 * hZ + 2]))
        # granular material (:399-446)
        rng = np.random.default_rng(self.m_seed)
        if self.m_in_layers:
            pts, z = [], r
            while z < self.m_init_depth:
                pts.append(self._sample((hX, 0, z), (hX - r, hY - r, 0), delta, rng))
                z += delta
            pts = np.concatenate(pts)
        else:
            pts = self._sample((hX, 0, D / 2), (hX - r, hY - r, D / 2 - r), delta, rng)
        self.m_num_particles = pts.shape[0]
        mass = self.m_rho_g * 4.0 / 3.0 * np.pi * self.m_radius_g ** 3
        self.m_particles_start = b.add_bodies(mass, (0.4 * mass * self.m_radius_g ** 2,) * 3, pts)
        ids = np.arange(self.m_particles_start, self.m_particles_start + self.m_num_particles)
        b.add_shapes(SPHERE, ids, np.full((ids.size, 1), self.m_radius_g))
        self.m_mass_g, self.m_inertia_g = mass, 0.4 * mass * self.m_radius_g ** 2
        settings = {k: (list(getattr(st, k)) if hasattr(getattr(st, k), "__len__") else getattr(st, k))
                    for k, _ in st._fields_ if k != "reserved"}
        load_scene(self.m_system, b.finish([self.m_material], settings))
        self.m_system.SetActiveBoundingBox(*self.m_active_box)
        self.m_tags = np.concatenate([[-1], TAG_PARTICLES + np.arange(self.m_num_particles)])
        if self.m_use_checkpoint:
            self._read_checkpoint(os.path.join(self.m_node_out_dir, self.m_checkpoint_filename))
        self.m_init_height = self.CalcCurrentHeight() + self.m_radius_g
        self.m_constructed = True

    # ---- settling (GOMP:588-698) ---------------------------------------------------------------------------------
    def Settle(self, check_every=50):
        """Simulate until the settling time is up or (threshold mode) the kinetic energy has dropped below it.  The
        reference looks at the energy after every step; here every `check_every` steps (a download per look)."""
        self.Construct()
        depth0, eta0 = self.CalculatePackingDensity()
        total_steps = int(np.ceil(self.m_time_settling / self.m_step_size))
        steps, time, max_contacts, cum_contacts, KE = 0, 0.0, 0, 0, 0.0
        while True:
            n = check_every if not self.m_fixed_settling_duration else min(check_every, total_steps - steps)
            self.m_system.DoSteps(max(n, 1))
            steps += max(n, 1)
            time = steps * self.m_step_size
            nc = self.m_system.GetNumContacts()
            cum_contacts += nc * max(n, 1)
            max_contacts = max(max_contacts, nc)
            if self.m_fixed_settling_duration:
                if steps >= total_steps:
                    KE = self.CalcTotalKineticEnergy()
                    break
            elif time > 0.1:
                KE = self.CalcTotalKineticEnergy()
                if KE <= self.m_KE_settling:
                    break
        self.m_init_height = self.CalcCurrentHeight() + self.m_radius_g
        depth1, eta1 = self.CalculatePackingDensity()
        stats = {"Number particles": self.m_num_particles, "Initial material depth": depth0,
                 "Initial packing density": eta0, "Final material depth": depth1, "Final packing density": eta1,
                 "Final kinetic energy": KE, "Average number of contacts": cum_contacts / steps,
                 "Maximum number contacts": max_contacts, "Settling duration": time}
        with open(os.path.join(self.m_node_out_dir, "settling_stats.info"), "w") as f:  # :680-693
            for k, v in stats.items():
                f.write(f"{k + ':':28s}{v}\n")
        return stats

    def _particle_state(self):
        self.m_system.Synchronize()
        return self.m_system.GetBodyState(self.m_particles_start, self.m_num_particles)

    def CalcTotalKineticEnergy(self):
        """GOMP:700-710: 1/2 sum (m |v|^2 + w . J w) over the particles (a sphere's J is isotropic: frame-free)."""
        _, _, v = self._particle_state()
        return 0.5 * float((self.m_mass_g * (v[:, :3] ** 2).sum() + self.m_inertia_g * (v[:, 3:] ** 2).sum()))

    def CalcCurrentHeight(self):
        """GOMP:712-719: z of the highest particle centre."""
        p, _, _ = self._particle_state()
        return float(p[:, 2].max())

    def CalculatePackingDensity(self):
        """GOMP:724-739 -> (depth, packing density)."""
        depth = self.CalcCurrentHeight()
        Vt = self.m_dimX * self.m_dimY * depth
        Vs = self.m_num_particles * 4.0 / 3.0 * np.pi * self.m_radius_g ** 3
        return depth, Vs / Vt

    def GetInitHeight(self):
        return self.m_init_height

    def GetNumContacts(self):
        return self.m_system.GetNumContacts()

    # ---- proxies (GOMP:749-858) ------------------------------------------------------------------------------------
    PROXY_FAMILY = (1 << 1, 0x7FFF & ~(1 << 1))  # SetFamily(1) + DisallowCollisionsWith(1)

    def OnInitialize(self):
        """Create the proxy bodies of every object added so far (ChVehicleCosimTerrainNodeChrono::OnInitialize ->
        CreateMeshProxy / CreateRigidProxy)."""
        self.Construct()
        for i, obj in enumerate(self.m_objects):
            if i in self.m_proxies:
                continue
            if obj["kind"] == "mesh":
                self.CreateMeshProxy(i)
            else:
                self.CreateRigidProxy(i)

    def CreateMeshProxy(self, i):
        """One body per mesh face with a triangle shape whose vertices are rewritten at every exchange (:749-790)."""
        obj, s = self.m_objects[i], self.m_system
        nt = obj["tris"].shape[0]
        mass_p = obj["mass"] / nt
        inertia_p = 1e-3 * mass_p * 0.1
        length = 0.1
        act = None if not self.m_fixed_proxies else np.zeros(nt, dtype=np.int8)
        first = s.AddBodies(np.full(nt, mass_p), np.full((nt, 3), inertia_p), np.zeros((nt, 3)), active=act)
        dummy = np.tile(np.array([[length, 0, 0], [0, length, 0], [0, 0, length]], float), (nt, 1))
        v0 = s.AddShapeVertices(dummy)
        dims = np.zeros((nt, 4))
        dims[:, 0], dims[:, 1] = v0 + 3 * np.arange(nt), 3
        s.AddShapes(np.full(nt, TRIANGLE), np.arange(first, first + nt), dims,
                    family=np.tile(np.asarray(self.PROXY_FAMILY, np.int16), (nt, 1)))
        self.m_proxies[i] = ProxyBodySet(first, nt, v0, obj["tris"])

    def CreateRigidProxy(self, i):
        """One body carrying the object's contact shapes (:792-822)."""
        obj, s = self.m_objects[i], self.m_system
        act = None if not self.m_fixed_proxies else np.zeros(1, dtype=np.int8)
        first = s.AddBodies([obj["mass"]], [[1.0, 1.0, 1.0]], [[0.0, 0.0, 0.0]], active=act)
        fam = np.asarray(self.PROXY_FAMILY, np.int16)
        for c, rad in obj["spheres"]:
            s.AddShapes([SPHERE], [first], [[rad, 0, 0, 0]], pos_local=[c], family=[fam])
        for c, h in obj["boxes"]:
            s.AddShapes([BOX], [first], [[h[0], h[1], h[2], 0]], pos_local=[c], family=[fam])
        self.m_proxies[i] = ProxyBodySet(first, 1)

    def UpdateMeshProxy(self, i, vpos, vvel):
        """MeshState -> proxies (:858-911): body at the face centroid with identity orientation, velocity = mean of the
        vertex velocities, zero angular velocity, triangle vertices relative to the centroid."""
        px = self.m_proxies[i]
        vpos, vvel = np.asarray(vpos, float).reshape(-1, 3), np.asarray(vvel, float).reshape(-1, 3)
        tri = px.idx_verts
        pA, pB, pC = vpos[tri[:, 0]], vpos[tri[:, 1]], vpos[tri[:, 2]]
        pos = (pA + pB + pC) * (1.0 / 3.0)
        vel = np.zeros((px.n, 6))
        vel[:, :3] = (vvel[tri[:, 0]] + vvel[tri[:, 1]] + vvel[tri[:, 2]]) * (1.0 / 3.0)
        rot = np.tile([1.0, 0.0, 0.0, 0.0], (px.n, 1))
        self.m_system.SetBodyState(px.first_body, pos, rot, vel)
        local = np.stack([pA - pos, pB - pos, pC - pos], 1).reshape(-1, 3)
        self.m_system.UpdateShapeVertices(px.first_vertex, local)

    def UpdateRigidProxy(self, i, pos, rot, lin_vel, ang_vel):
        """BodyState -> proxy (:913-919); ang_vel is in the parent frame (SetAngVelParent)."""
        px = self.m_proxies[i]
        q = np.asarray(rot, float)
        w = np.asarray(ang_vel, float)
        # local angular velocity: q^-1 w q
        e0, ev = q[0], q[1:]
        t = 2.0 * np.cross(ev, w)
        w_loc = w - e0 * t + np.cross(ev, t)
        self.m_system.SetBodyState(px.first_body, [pos], [q], [np.concatenate([lin_vel, w_loc])])

    def GetForceMeshProxy(self, i):
        """Proxies -> MeshContact (:946-1006): a face's contact force goes in thirds to its vertices; returns the
        indices of the vertices in contact (ascending) and their forces."""
        px = self.m_proxies[i]
        f, _ = self.m_system.GetBodyContactForces()
        f = f[px.first_body:px.first_body + px.n]
        hit = np.any(f != 0.0, axis=1)
        nv = int(px.idx_verts.max()) + 1
        acc = np.zeros((nv, 3))
        touched = np.zeros(nv, dtype=bool)
        third = f[hit] * (1.0 / 3.0)
        for k in range(3):
            np.add.at(acc, px.idx_verts[hit, k], third)
            touched[px.idx_verts[hit, k]] = True
        vidx = np.nonzero(touched)[0]
        return vidx, acc[vidx]

    def GetForceRigidProxy(self, i):
        """Proxy -> TerrainForce (:1009-1015): (point, force, moment); the point is the body origin."""
        px = self.m_proxies[i]
        f, t = self.m_system.GetBodyContactForces()
        return np.zeros(3), f[px.first_body].copy(), t[px.first_body].copy()

    def OnAdvance(self, step_size):
        """ChVehicleCosimTerrainNodeChrono::OnAdvance + CalculateContactForces (:1018-1024): whole solver steps of
        m_step_size until step_size is covered."""
        n = max(int(round(step_size / self.m_step_size)), 1)
        self.m_system.DoSteps(n)
        self.m_system.Synchronize()

    # ---- output (GOMP:1043-1084) ---------------------------------------------------------------------------------
    @staticmethod
    def _rot_dt(q, w_loc):
        """ChBody::GetRotDt: 1/2 q * (0, w_local)."""
        e0, e1, e2, e3 = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        wx, wy, wz = w_loc[:, 0], w_loc[:, 1], w_loc[:, 2]
        return 0.5 * np.stack([-e1 * wx - e2 * wy - e3 * wz, e0 * wx + e2 * wz - e3 * wy,
                               e0 * wy + e3 * wx - e1 * wz, e0 * wz + e1 * wy - e2 * wx], 1)

    def WriteParticleInformation(self, filename):
        p, _, v = self._particle_state()
        np.savetxt(filename, np.concatenate([p, v[:, :3]], 1), fmt="%.6g")

    def WriteCheckpoint(self, filename, precision=6):
        """Time, particle count, then `tag pos rot pos_dt rot_dt` per particle, space separated, one trailing space per
        item as ChWriterCSV leaves it (:1063-1083).  precision 6 is the reference's (default stream format); 17 keeps
        every bit."""
        p, q, v = self._particle_state()
        rdt = self._rot_dt(q, v[:, 3:])
        fmt = "{:." + str(int(precision)) + "g}"
        path = os.path.join(self.m_node_out_dir, filename)
        with open(path, "w") as f:
            f.write(fmt.format(self.m_system.ch_time) + " \n")
            f.write(f"{self.m_num_particles} \n")
            rows = np.concatenate([p, q, v[:, :3], rdt], 1)
            for tag, row in zip(self.m_tags[1:], rows):
                f.write(f"{tag} " + " ".join(fmt.format(x) for x in row) + " \n")
        return path

    def _read_checkpoint(self, path):
        """GOMP:452-497: overwrite the particle states; the particle count must match."""
        with open(path) as f:
            f.readline()  # time: read and discarded
            n = int(f.readline().split()[0])
            if n != self.m_num_particles:
                raise ValueError("inconsistent number of particles in checkpoint file")
            rows = np.loadtxt(f, ndmin=2)
        if rows.shape != (n, 15) or not np.array_equal(rows[:, 0].astype(np.int64), self.m_tags[1:]):
            raise ValueError("checkpoint rows do not match the particles of this node")
        pos, q, rdt = rows[:, 1:4], rows[:, 4:8], rows[:, 11:15]
        # w_local = 2 q^-1 * rot_dt (vector part)
        e0, e1, e2, e3 = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        r0, r1, r2, r3 = rdt[:, 0], rdt[:, 1], rdt[:, 2], rdt[:, 3]
        w = 2.0 * np.stack([e0 * r1 - e1 * r0 - e2 * r3 + e3 * r2, e0 * r2 + e1 * r3 - e2 * r0 - e3 * r1,
                            e0 * r3 - e1 * r2 + e2 * r1 - e3 * r0], 1)
        self.m_system.SetBodyState(self.m_particles_start, pos, q, np.concatenate([rows[:, 8:11], w], 1))
