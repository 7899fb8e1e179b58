"""Turns a morphology table (robots/mjcf_tables.py) into the link-per-DoF articulation model the
K2 kernel consumes (``OigeModelDesc`` of include/oige_b200.h).

Conventions
  * link 0 = floating root (the MJCF root body).  Every hinge joint becomes one link; a body with
    k joints becomes a chain of k links of which only the last carries the body's mass, geoms and
    (for feet) the force sensor -- the intermediate links are massless.
  * every link frame is oriented like the MJCF frame of the body it belongs to (at q = 0) and has
    its origin at the joint anchor.
  * DoFs are numbered breadth-first over BODIES (docs/transfering_policies_from_isaac_gym.md:39-54),
    inside a body in the reference's listed order (tasks/humanoid.py:82-107).
  * capsules contribute two end-cap collision spheres, spheres one.
"""
import math
from dataclasses import dataclass, field
from typing import Dict, List

import numpy as np

from omniisaacgymenvs_b200.robots import mjcf_tables


def _quat_to_mat(q):
    w, x, y, z = q
    n = math.sqrt(w * w + x * x + y * y + z * z)
    w, x, y, z = w / n, x / n, y / n, z / n
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)],
    ])


def _geom_inertial(geom, density):
    """(mass, com[3], inertia-about-com 3x3) of a capsule / sphere / box of uniform density."""
    if geom[0] == "box":  # ("box", centre, full sizes), axis-aligned in the body frame
        _, pos, size = geom
        sx, sy, sz = (float(v) for v in size)
        m = density * sx * sy * sz
        return m, np.array(pos, float), np.diag([m * (sy * sy + sz * sz) / 12.0, m * (sx * sx + sz * sz) / 12.0,
                                                 m * (sx * sx + sy * sy) / 12.0])
    if geom[0] == "sphere":
        _, pos, r = geom
        m = density * 4.0 / 3.0 * math.pi * r ** 3
        return m, np.array(pos, float), np.eye(3) * (0.4 * m * r * r)
    _, a, b, r = geom
    a, b = np.array(a, float), np.array(b, float)
    length = np.linalg.norm(b - a)
    h = 0.5 * length
    m_cyl = density * math.pi * r * r * length
    m_hs = density * 2.0 / 3.0 * math.pi * r ** 3
    m = m_cyl + 2 * m_hs
    i_ax = 0.5 * m_cyl * r * r + 2 * m_hs * 0.4 * r * r
    i_tr = m_cyl * (length * length / 12.0 + r * r / 4.0) + 2 * (
        m_hs * (83.0 / 320.0) * r * r + m_hs * (h + 3.0 * r / 8.0) ** 2
    )
    z = (b - a) / length if length > 0 else np.array([0.0, 0.0, 1.0])
    inertia = i_tr * (np.eye(3) - np.outer(z, z)) + i_ax * np.outer(z, z)
    return m, 0.5 * (a + b), inertia


@dataclass
class ArticulationModel:
    name: str
    num_links: int
    num_dof: int
    parent: np.ndarray            # [L] int
    dof_index: np.ndarray         # [L] int (-1 root)
    joint_type: np.ndarray        # [L] int: 0 hinge, 1 prismatic
    joint_pos: np.ndarray         # [L,3]
    joint_quat: np.ndarray        # [L,4] wxyz
    joint_axis: np.ndarray        # [L,3]
    mass: np.ndarray              # [L]
    com: np.ndarray               # [L,3]
    inertia: np.ndarray           # [L,6] xx yy zz xy xz yz
    armature: np.ndarray
    damping: np.ndarray
    stiffness: np.ndarray
    spring_ref: np.ndarray
    limit_lower: np.ndarray       # [L] rad
    limit_upper: np.ndarray
    drive_stiffness: np.ndarray   # [L] PD position-drive gains (0 = no drive)
    drive_damping: np.ndarray
    drive_max_force: np.ndarray   # <= 0: unlimited
    sphere_link: np.ndarray       # [K]
    sphere_pos: np.ndarray        # [K,3]
    sphere_radius: np.ndarray     # [K]
    sensor_link: np.ndarray       # [S]
    dof_names: List[str]
    dof_body_names: List[str]
    joint_gears: np.ndarray       # [D] in DoF order
    spawn_height: float
    link_names: List[str] = field(default_factory=list)
    # physics constants (cfg/task/*.yaml sim block + documented substitutions)
    gravity: tuple = (0.0, 0.0, -9.81)
    contact_stiffness: float = 2.0e4
    contact_damping: float = 5.0e2
    friction_mu: float = 1.0
    friction_viscous: float = 2.0e3
    limit_stiffness: float = 2.0e3
    limit_damping: float = 2.0e1
    # Isaac Sim USD defaults the reference's assets run with (docs/transfering_policies_from_isaac_gym.md:71-77,85-87):
    # angular damping 0.05 on every body, max angular velocity 5729.58 deg/s, max joint velocity 1e6 deg/s.  (The Isaac
    # Gym values of the same table -- 0.0, 64 rad/s, 100 rad/s -- are a `sim.b200` override away.)
    max_joint_velocity: float = 1.0e6 * math.pi / 180.0
    angular_damping: float = 0.05
    max_angular_velocity: float = 5729.58008 * math.pi / 180.0
    max_depenetration_velocity: float = 10.0   # cfg/task/Humanoid.yaml:62,90
    self_contact_stiffness: float = 1.0e4      # explicit capsule-capsule penalty (DESIGN.md "Self-collision")
    self_contact_damping: float = 5.0e1
    enable_self_collisions: bool = False       # cfg/task/Humanoid.yaml:80 (True), Ant.yaml:82 (False)
    collision_pairs: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), np.int32))  # [P,2] link pairs
    # self-collision shape of a link: one capsule (segment a-b in the link frame, radius; radius 0 = none)
    capsule_a: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    capsule_b: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    capsule_radius: np.ndarray = field(default_factory=lambda: np.zeros(0))
    fixed_base: bool = False      # link 0 welded to the world

    @property
    def num_sensors(self) -> int:
        return int(len(self.sensor_link))

    def dof_limits(self) -> np.ndarray:
        """[D,2] (lower, upper) in DoF order -- what ArticulationView.get_dof_limits()[0] returns."""
        out = np.zeros((self.num_dof, 2), np.float32)
        for link in range(1, self.num_links):
            out[self.dof_index[link], 0] = self.limit_lower[link]
            out[self.dof_index[link], 1] = self.limit_upper[link]
        return out

    def total_mass(self) -> float:
        return float(self.mass.sum())


def _gear_for(table, joint_name: str) -> float:
    for key, val in table["gears"].items():
        if joint_name.endswith(key) or joint_name.startswith(key):
            return val
    raise KeyError(joint_name)


def build_model(name: str, **overrides) -> ArticulationModel:
    table = mjcf_tables.ROBOTS[name]
    bodies = table["bodies"]()
    by_name = {b["name"]: b for b in bodies}
    # body levels for the breadth-first DoF numbering
    level: Dict[str, int] = {}
    for b in bodies:
        level[b["name"]] = 0 if b["parent"] is None else level[b["parent"]] + 1
    bfs_bodies = sorted(bodies, key=lambda b: (level[b["name"]], bodies.index(b)))
    dof_of_joint: Dict[str, int] = {}
    dof_names: List[str] = []
    dof_body_names: List[str] = []
    for b in bfs_bodies:
        order = table["dof_order_in_body"].get(b["name"], [j["name"] for j in b["joints"]])
        for jn in order:
            dof_of_joint[jn] = len(dof_names)
            dof_names.append(jn)
            dof_body_names.append(b["name"])

    parent, dof_index, jpos, jquat, jaxis, jtype = [-1], [-1], [(0, 0, 0)], [(1, 0, 0, 0)], [(0, 0, 1)], [0]
    arm, damp, stiff, lo, hi, link_names = [0.0], [0.0], [0.0], [0.0], [0.0], [bodies[0]["name"]]
    last_link: Dict[str, int] = {bodies[0]["name"]: 0}
    last_anchor: Dict[str, np.ndarray] = {bodies[0]["name"]: np.zeros(3)}
    for b in bodies[1:]:  # document order guarantees parents come first
        if not b["joints"]:
            raise ValueError(f"body {b['name']} has no joint: fixed children are not supported")
        pl = last_link[b["parent"]]
        pa = last_anchor[b["parent"]]
        Rb = _quat_to_mat(b["quat"])
        prev_anchor = None
        for k, j in enumerate(b["joints"]):
            anchor = np.array(j["anchor"], float)
            if k == 0:
                jp = -pa + np.array(b["pos"], float) + Rb @ anchor
                jq = tuple(float(v) for v in np.array(b["quat"], float) / np.linalg.norm(b["quat"]))
                par = pl
            else:
                jp = anchor - prev_anchor
                jq = (1.0, 0.0, 0.0, 0.0)
                par = len(parent) - 1
            ax = np.array(j["axis"], float)
            ax = ax / np.linalg.norm(ax)
            parent.append(par)
            dof_index.append(dof_of_joint[j["name"]])
            jpos.append(tuple(jp))
            jquat.append(jq)
            jaxis.append(tuple(ax))
            arm.append(j["armature"]); damp.append(j["damping"]); stiff.append(j["stiffness"])
            slide = j.get("type", "hinge") == "slide"
            jtype.append(1 if slide else 0)
            # hinge ranges are in degrees, slide ranges in metres; None = unlimited (URDF "continuous")
            if j["range"] is None:
                lo.append(-1.0e30); hi.append(1.0e30)
            elif slide:
                lo.append(float(j["range"][0])); hi.append(float(j["range"][1]))
            else:
                lo.append(math.radians(j["range"][0])); hi.append(math.radians(j["range"][1]))
            link_names.append(j["name"])
            prev_anchor = anchor
        last_link[b["name"]] = len(parent) - 1
        last_anchor[b["name"]] = prev_anchor

    L = len(parent)
    mass = np.zeros(L); com = np.zeros((L, 3)); inertia = np.zeros((L, 6))
    sph_link, sph_pos, sph_r = [], [], []
    for b in bodies:
        link = last_link[b["name"]]
        off = last_anchor[b["name"]]
        parts = [_geom_inertial(g, table["density"]) for g in b["geoms"] if g[0] != "csphere"]
        if b.get("mass") is not None:  # URDF-style explicit body mass: rescale the uniform-density parts
            k = float(b["mass"]) / sum(p[0] for p in parts)
            parts = [(p[0] * k, p[1], p[2] * k) for p in parts]
        m = sum(p[0] for p in parts)
        c = sum(p[0] * p[1] for p in parts) / m
        I = np.zeros((3, 3))
        for pm, pc, pi in parts:
            d = pc - c
            I += pi + pm * (np.dot(d, d) * np.eye(3) - np.outer(d, d))
        mass[link] = m
        com[link] = c - off
        inertia[link] = [I[0, 0], I[1, 1], I[2, 2], I[0, 1], I[0, 2], I[1, 2]]
        for g in b["geoms"]:
            if g[0] == "box":
                continue  # boxes carry mass only
            if g[0] == "csphere":  # collision-only sphere (no mass)
                sph_link.append(link); sph_pos.append(np.array(g[1], float) - off); sph_r.append(g[2])
                continue
            if g[0] == "sphere":
                sph_link.append(link); sph_pos.append(np.array(g[1], float) - off); sph_r.append(g[2])
            else:
                for end in (g[1], g[2]):
                    sph_link.append(link); sph_pos.append(np.array(end, float) - off); sph_r.append(g[3])
    gears = np.zeros(len(dof_names), np.float32)
    for jn, d in dof_of_joint.items():
        gears[d] = _gear_for(table, jn)
    model = ArticulationModel(
        name=name, num_links=L, num_dof=L - 1,
        parent=np.array(parent, np.int32), dof_index=np.array(dof_index, np.int32),
        joint_type=np.array(jtype, np.int32), fixed_base=bool(table.get("fixed_base", False)),
        joint_pos=np.array(jpos, np.float64), joint_quat=np.array(jquat, np.float64),
        joint_axis=np.array(jaxis, np.float64), mass=mass, com=com, inertia=inertia,
        armature=np.array(arm), damping=np.array(damp), stiffness=np.array(stiff), spring_ref=np.zeros(L),
        limit_lower=np.array(lo), limit_upper=np.array(hi),
        drive_stiffness=np.zeros(L), drive_damping=np.zeros(L), drive_max_force=np.zeros(L),
        sphere_link=np.array(sph_link, np.int32), sphere_pos=np.array(sph_pos, np.float64).reshape(-1, 3),
        sphere_radius=np.array(sph_r, np.float64),
        sensor_link=np.array([last_link[s] for s in table["sensor_bodies"]], np.int32),
        dof_names=dof_names, dof_body_names=dof_body_names, joint_gears=gears,
        spawn_height=table["spawn_height"], link_names=link_names,
    )
    # self-collision candidates, PhysX articulation semantics: every pair of BODIES with collision shapes except a body
    # and its parent (jointed links never collide); expressed on the links that carry the bodies' spheres
    body_parent = {b["name"]: b["parent"] for b in bodies}
    with_spheres = [b["name"] for b in bodies if any(int(l) == last_link[b["name"]] for l in sph_link)]
    pairs = []
    for ia, a in enumerate(with_spheres):
        for b in with_spheres[ia + 1:]:
            if body_parent[a] == b or body_parent[b] == a:
                continue
            pairs.append((last_link[a], last_link[b]))
    model.collision_pairs = np.array(pairs, np.int32).reshape(-1, 2)
    # one self-collision capsule per body: the table's `sc_capsule` proxy, else the body's single capsule / sphere geom
    model.capsule_a, model.capsule_b, model.capsule_radius = np.zeros((L, 3)), np.zeros((L, 3)), np.zeros(L)
    for b in bodies:
        link, off = last_link[b["name"]], last_anchor[b["name"]]
        shape = b.get("sc_capsule")
        solid = [g for g in b["geoms"] if g[0] in ("capsule", "sphere")]
        if shape is None and len(solid) == 1:
            g = solid[0]
            shape = (g[1], g[2], g[3]) if g[0] == "capsule" else (g[1], g[1], g[2])
        if shape is None:
            continue  # (a multi-geom body without a proxy takes no part in self-collision)
        model.capsule_a[link] = np.array(shape[0], float) - off
        model.capsule_b[link] = np.array(shape[1], float) - off
        model.capsule_radius[link] = float(shape[2])
    model.collision_pairs = np.array([pr for pr in model.collision_pairs
                                      if model.capsule_radius[pr[0]] > 0 and model.capsule_radius[pr[1]] > 0],
                                     np.int32).reshape(-1, 2)
    model.enable_self_collisions = bool(table.get("self_collision", False))
    drive = table.get("drive")
    if drive:  # PD position drive on every joint (tasks/anymal.py:128)
        model.drive_stiffness[1:] = drive["stiffness"]
        model.drive_damping[1:] = drive["damping"]
        model.drive_max_force[1:] = drive["max_force"]
    for k, v in overrides.items():
        if not hasattr(model, k):
            raise AttributeError(k)
        setattr(model, k, v)
    return model


def fill_desc(model: ArticulationModel, desc) -> None:
    """Fills a ctypes ``OigeModelDesc`` (or the oracle's identical struct) from the model."""
    L, K, S = model.num_links, len(model.sphere_link), model.num_sensors
    desc.num_links, desc.num_spheres, desc.num_sensors = L, K, S
    desc.fixed_base = 1 if model.fixed_base else 0
    for i in range(L):
        desc.parent[i] = int(model.parent[i])
        desc.dof_index[i] = int(model.dof_index[i])
        desc.joint_type[i] = int(model.joint_type[i])
        for k in range(3):
            desc.joint_pos[i][k] = float(model.joint_pos[i, k])
            desc.joint_axis[i][k] = float(model.joint_axis[i, k])
            desc.com[i][k] = float(model.com[i, k])
        for k in range(4):
            desc.joint_quat[i][k] = float(model.joint_quat[i, k])
        for k in range(6):
            desc.inertia[i][k] = float(model.inertia[i, k])
        desc.mass[i] = float(model.mass[i])
        desc.armature[i] = float(model.armature[i])
        desc.damping[i] = float(model.damping[i])
        desc.stiffness[i] = float(model.stiffness[i])
        desc.spring_ref[i] = float(model.spring_ref[i])
        desc.limit_lower[i] = float(model.limit_lower[i])
        desc.limit_upper[i] = float(model.limit_upper[i])
        desc.drive_stiffness[i] = float(model.drive_stiffness[i])
        desc.drive_damping[i] = float(model.drive_damping[i])
        desc.drive_max_force[i] = float(model.drive_max_force[i])
    for s in range(K):
        desc.sphere_link[s] = int(model.sphere_link[s])
        desc.sphere_radius[s] = float(model.sphere_radius[s])
        for k in range(3):
            desc.sphere_pos[s][k] = float(model.sphere_pos[s, k])
    for s in range(S):
        desc.sensor_link[s] = int(model.sensor_link[s])
    for k in range(3):
        desc.gravity[k] = float(model.gravity[k])
    desc.contact_stiffness = model.contact_stiffness
    desc.contact_damping = model.contact_damping
    desc.friction_mu = model.friction_mu
    desc.friction_viscous = model.friction_viscous
    desc.limit_stiffness = model.limit_stiffness
    desc.limit_damping = model.limit_damping
    desc.max_joint_velocity = model.max_joint_velocity
    desc.angular_damping = model.angular_damping
    desc.max_angular_velocity = model.max_angular_velocity
    desc.max_depenetration_velocity = model.max_depenetration_velocity
    desc.self_contact_stiffness = model.self_contact_stiffness
    desc.self_contact_damping = model.self_contact_damping
    pairs = model.collision_pairs if model.enable_self_collisions else model.collision_pairs[:0]
    if len(pairs) > len(desc.pair_link_a):
        raise ValueError(f"{len(pairs)} self-collision link pairs exceed OIGE_MAX_PAIRS")
    desc.num_collision_pairs = len(pairs)
    for k, (a, b) in enumerate(pairs):
        desc.pair_link_a[k], desc.pair_link_b[k] = int(a), int(b)
    for i in range(L):
        has = len(model.capsule_radius) == L and model.capsule_radius[i] > 0
        desc.capsule_radius[i] = float(model.capsule_radius[i]) if has else 0.0
        for k in range(3):
            desc.capsule_a[i][k] = float(model.capsule_a[i, k]) if has else 0.0
            desc.capsule_b[i][k] = float(model.capsule_b[i, k]) if has else 0.0
