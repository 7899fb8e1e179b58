"""Morphology tables of the two locomotion robots.

The reference does not contain the robots: it references ``humanoid_instanceable.usd`` /
``ant_instanceable.usd`` on a Nucleus server (robots/articulations/humanoid.py:57, ant.py:57),
which are USD conversions of the public ``nv_humanoid.xml`` / ``nv_ant.xml`` MJCF files of
IsaacGymEnvs.  The tables below restate those public MJCF models (bodies, hinge joints, capsule /
sphere geoms) as recalled -- PROVENANCE: not pinned by anything under /root/reference.  What the
reference DOES pin is cross-checked in tests/test_config_models.py: DoF counts (21 / 8), the breadth-
first joint order with its body names and gears (tasks/humanoid.py:82-107, tasks/ant.py:80),
sensor counts (2 / 4, from the observation sizes 87 / 60) and spawn heights (1.34 / 0.5).

Units: metres, degrees for joint ranges (converted to radians by the builder), MuJoCo default
geom density 1000 kg/m^3 for the humanoid, 5 kg/m^3 for the ant (as in nv_ant.xml).
"""

# body: name, parent, pos, quat(wxyz), joints[(name, anchor, axis, range_deg, stiffness, damping, armature)],
#       geoms[("capsule", from, to, radius) | ("sphere", pos, radius)]

_HJ = dict(stiffness=5.0, damping=0.1, armature=0.007)  # humanoid joint class defaults


def _hj(name, anchor, axis, rng, **kw):
    d = dict(_HJ)
    d.update(kw)
    return dict(name=name, anchor=anchor, axis=axis, range=rng, **d)


def _leg(side: str):
    s = -1.0 if side == "right" else 1.0  # right leg sits at y = -0.1
    thigh = dict(
        name=f"{side}_thigh", parent="pelvis", pos=(0.0, s * 0.1, -0.04), quat=(1, 0, 0, 0),
        joints=[
            _hj(f"{side}_hip_x", (0, 0, 0), (1, 0, 0), (-45, 15) if side == "right" else (-15, 45),
                stiffness=10.0, damping=5.0, armature=0.01),
            _hj(f"{side}_hip_z", (0, 0, 0), (0, 0, 1), (-60, 35) if side == "right" else (-35, 60),
                stiffness=10.0, damping=5.0, armature=0.01),
            _hj(f"{side}_hip_y", (0, 0, 0), (0, 1, 0), (-120, 45), stiffness=20.0, damping=5.0, armature=0.01),
        ],
        geoms=[("capsule", (0, 0, 0), (0, -s * 0.01, -0.34), 0.06)],
    )
    shin = dict(
        name=f"{side}_shin", parent=f"{side}_thigh", pos=(0.0, -s * 0.01, -0.403), quat=(1, 0, 0, 0),
        joints=[_hj(f"{side}_knee", (0, 0, 0.02), (0, -1, 0), (-160, 2))],
        geoms=[("capsule", (0, 0, 0), (0, 0, -0.3), 0.049)],
    )
    foot = dict(
        name=f"{side}_foot", parent=f"{side}_shin", pos=(0.0, 0.0, -0.39), quat=(1, 0, 0, 0),
        joints=[
            _hj(f"{side}_ankle_y", (0, 0, 0.08), (0, 1, 0), (-50, 50), stiffness=6.0),
            _hj(f"{side}_ankle_x", (0, 0, 0.04), (1, 0, 0.5), (-50, 50), stiffness=3.0),
        ],
        geoms=[
            ("capsule", (-0.07, -0.02, 0), (0.14, -0.04, 0), 0.027),
            ("capsule", (-0.07, 0.0, 0), (0.14, 0.02, 0), 0.027),
        ] if side == "right" else [
            ("capsule", (-0.07, 0.02, 0), (0.14, 0.04, 0), 0.027),
            ("capsule", (-0.07, 0.0, 0), (0.14, -0.02, 0), 0.027),
        ],
        # self-collision proxy: one capsule along the foot enclosing the two sole capsules
        sc_capsule=((-0.07, -s * 0.01, 0), (0.14, -s * 0.01, 0), 0.045),
    )
    return thigh, shin, foot


def _arm(side: str):
    s = -1.0 if side == "right" else 1.0
    upper = dict(
        name=f"{side}_upper_arm", parent="torso", pos=(0.0, s * 0.17, 0.06), quat=(1, 0, 0, 0),
        joints=[
            _hj(f"{side}_shoulder1", (0, 0, 0), (2, -s * 1, 1), (-90, 70) if side == "right" else (-70, 90)),
            _hj(f"{side}_shoulder2", (0, 0, 0), (0, s * 1, 1), (-90, 70) if side == "right" else (-70, 90)),
        ],
        geoms=[("capsule", (0, 0, 0), (0.16, s * 0.16, -0.16), 0.04)],
    )
    lower = dict(
        name=f"{side}_lower_arm", parent=f"{side}_upper_arm", pos=(0.18, s * 0.18, -0.18), quat=(1, 0, 0, 0),
        joints=[_hj(f"{side}_elbow", (0, 0, 0), (0, -1, -s * 1), (-90, 50), stiffness=2.0)],
        geoms=[
            ("capsule", (0.01, -s * 0.01, 0.01), (0.17, -s * 0.17, 0.17), 0.031),
            ("sphere", (0.18, -s * 0.18, 0.18), 0.04),  # hand
        ],
        sc_capsule=((0.01, -s * 0.01, 0.01), (0.18, -s * 0.18, 0.18), 0.04),  # forearm + hand
    )
    return upper, lower


def humanoid_bodies():
    rt, rs, rf = _leg("right")
    lt, ls, lf = _leg("left")
    rua, rla = _arm("right")
    lua, lla = _arm("left")
    torso = dict(
        name="torso", parent=None, pos=(0, 0, 0), quat=(1, 0, 0, 0), joints=[],
        geoms=[
            ("capsule", (0, -0.07, 0), (0, 0.07, 0), 0.07),
            ("capsule", (-0.01, -0.06, -0.12), (-0.01, 0.06, -0.12), 0.06),  # upper_waist
            ("sphere", (0, 0, 0.19), 0.09),  # head
        ],
        sc_capsule=((-0.01, 0, -0.12), (0, 0, 0.19), 0.11),  # upper waist .. head column
    )
    lower_waist = dict(
        name="lower_waist", parent="torso", pos=(-0.01, 0, -0.26), quat=(1.0, 0, -0.002, 0),
        joints=[
            _hj("abdomen_z", (0, 0, 0.065), (0, 0, 1), (-45, 45), stiffness=20.0, damping=5.0, armature=0.02),
            _hj("abdomen_y", (0, 0, 0.065), (0, 1, 0), (-75, 30), stiffness=20.0, damping=5.0, armature=0.01),
        ],
        geoms=[("capsule", (0, -0.06, 0), (0, 0.06, 0), 0.06)],
    )
    pelvis = dict(
        name="pelvis", parent="lower_waist", pos=(0, 0, -0.165), quat=(1.0, 0, -0.002, 0),
        joints=[_hj("abdomen_x", (0, 0, 0.1), (1, 0, 0), (-35, 35), stiffness=10.0, damping=5.0, armature=0.01)],
        geoms=[("capsule", (-0.02, -0.07, 0), (-0.02, 0.07, 0), 0.09)],  # butt
    )
    # order = MJCF document order; the builder derives the breadth-first DoF order from it
    return [torso, lower_waist, rua, lua, pelvis, rla, lla, rt, lt, rs, ls, rf, lf]


HUMANOID = dict(
    name="Humanoid",
    bodies=humanoid_bodies,
    density=1000.0,
    spawn_height=1.34,                    # tasks/humanoid.py:61
    sensor_bodies=["right_foot", "left_foot"],
    self_collision=True,                  # cfg/task/Humanoid.yaml:80 (enable_self_collisions: True; Ant.yaml:82 False)
    # per-DoF order inside a multi-DoF body as the reference lists it (tasks/humanoid.py:82-107:
    # thigh x (45), y (135), z (45)); kinematic chain order stays the MJCF one (x, z, y)
    dof_order_in_body={"right_thigh": ["right_hip_x", "right_hip_y", "right_hip_z"],
                       "left_thigh": ["left_hip_x", "left_hip_y", "left_hip_z"]},
    gears={"abdomen_z": 67.5, "abdomen_y": 67.5, "abdomen_x": 67.5, "shoulder1": 67.5, "shoulder2": 67.5,
           "elbow": 45.0, "hip_x": 45.0, "hip_y": 135.0, "hip_z": 45.0, "knee": 90.0, "ankle_y": 22.5,
           "ankle_x": 22.5},
)


_AJ = dict(stiffness=0.0, damping=0.1, armature=0.01)  # nv_ant.xml joint defaults


def ant_bodies():
    torso = dict(
        name="torso", parent=None, pos=(0, 0, 0), quat=(1, 0, 0, 0), joints=[],
        geoms=[
            ("sphere", (0, 0, 0), 0.25),
            ("capsule", (0, 0, 0), (0.2, 0.2, 0), 0.08),
            ("capsule", (0, 0, 0), (-0.2, 0.2, 0), 0.08),
            ("capsule", (0, 0, 0), (-0.2, -0.2, 0), 0.08),
            ("capsule", (0, 0, 0), (0.2, -0.2, 0), 0.08),
        ],
    )
    legs = []
    feet = []
    spec = [
        ("front_left", (0.2, 0.2), (-1, 1, 0), (30, 100)),
        ("front_right", (-0.2, 0.2), (1, 1, 0), (-100, -30)),
        ("left_back", (-0.2, -0.2), (-1, 1, 0), (-100, -30)),
        ("right_back", (0.2, -0.2), (1, 1, 0), (30, 100)),
    ]
    for i, (nm, (x, y), ankle_axis, ankle_rng) in enumerate(spec, start=1):
        legs.append(dict(
            name=f"{nm}_leg", parent="torso", pos=(x, y, 0), quat=(1, 0, 0, 0),
            joints=[dict(name=f"hip_{i}", anchor=(0, 0, 0), axis=(0, 0, 1), range=(-40, 40), **_AJ)],
            geoms=[("capsule", (0, 0, 0), (x, y, 0), 0.08)],
        ))
        feet.append(dict(
            name=f"{nm}_foot", parent=f"{nm}_leg", pos=(x, y, 0), quat=(1, 0, 0, 0),
            joints=[dict(name=f"ankle_{i}", anchor=(0, 0, 0), axis=ankle_axis, range=ankle_rng, **_AJ)],
            geoms=[("capsule", (0, 0, 0), (2 * x, 2 * y, 0), 0.08)],
        ))
    return [torso] + legs + feet


ANT = dict(
    name="Ant",
    bodies=ant_bodies,
    density=5.0,
    spawn_height=0.5,                     # tasks/ant.py:60
    sensor_bodies=["front_left_foot", "front_right_foot", "left_back_foot", "right_back_foot"],
    dof_order_in_body={},
    gears={"hip": 15.0, "ankle": 15.0},   # tasks/ant.py:80
)

def cartpole_bodies():
    """The Cartpole of ``/Isaac/Robots/Cartpole/cartpole.usd`` (robots/articulations/cartpole.py:52), a USD
    conversion of the public IsaacGymEnvs ``cartpole.urdf``, restated as recalled (PROVENANCE: unpinned;
    the reference pins only the joint names, tasks/cartpole.py:133-134, and the spawn height 2.0,
    tasks/cartpole.py:57): an 8 m rail welded to the world, a 1 kg cart sliding along y within +-4 m
    and a 1 kg, 1 m pole on a continuous hinge about x, its box centred 0.47 m above the hinge."""
    rail = dict(name="rail", parent=None, pos=(0, 0, 0), quat=(1, 0, 0, 0), joints=[],
                geoms=[("box", (0, 0, 0), (0.03, 8.0, 0.03))], mass=1.0)
    cart = dict(
        name="cart", parent="rail", pos=(0, 0, 0), quat=(1, 0, 0, 0), mass=1.0,
        joints=[dict(name="cartJoint", type="slide", anchor=(0, 0, 0), axis=(0, 1, 0), range=(-4.0, 4.0),
                     stiffness=0.0, damping=0.0, armature=0.0)],
        geoms=[("box", (0, 0, 0), (0.2, 0.25, 0.2))],
    )
    pole = dict(
        name="pole", parent="cart", pos=(0.12, 0, 0), quat=(1, 0, 0, 0), mass=1.0,
        joints=[dict(name="poleJoint", type="hinge", anchor=(0, 0, 0), axis=(1, 0, 0), range=None,
                     stiffness=0.0, damping=0.0, armature=0.0)],
        geoms=[("box", (0, 0, 0.47), (0.04, 0.06, 1.0))],
    )
    return [rail, cart, pole]


CARTPOLE = dict(
    name="Cartpole",
    bodies=cartpole_bodies,
    density=1000.0,                       # overridden per body by the explicit masses
    spawn_height=2.0,                     # tasks/cartpole.py:57
    fixed_base=True,                      # the rail is jointed to the world in the USD
    sensor_bodies=[],
    dof_order_in_body={},
    gears={"cartJoint": 1.0, "poleJoint": 1.0},  # efforts are formed by the task (tasks/cartpole.py:104-105)
)

def anymal_bodies():
    """ANYmal C (``/Isaac/Robots/ANYbotics/anymal_instanceable.usd``, robots/articulations/anymal.py:58), restated
    from the public ``anymal_c`` URDF as recalled (PROVENANCE: unpinned; the reference pins the 12 DoF names and
    their order, robots/articulations/anymal.py:72-83, the spawn height 0.62, tasks/anymal.py:100, the fall
    threshold 0.51 m, tasks/anymal.py:284, and the link names ``base`` / ``*_SHANK``, views/anymal_view.py:49-50).
    Per leg: HAA about x at (+-0.2999, +-0.104, 0) on the base, HFE about y at (+-0.0599, +-0.08381, 0), KFE about
    y at (0, +-0.1003, -0.285), foot (sphere r = 0.03) at (+-0.08795, +-0.01305, -0.33797) in the shank frame.
    Masses: base + attached boxes 25.2, hip 2.781, thigh 3.071, shank + foot 0.838 kg (52 kg in total).  Shapes
    only distribute those masses and place collision spheres (feet, knees, base corners)."""
    import math

    base = dict(
        name="base", parent=None, pos=(0, 0, 0), quat=(1, 0, 0, 0), joints=[], mass=25.2,
        geoms=[("box", (0, 0, 0.02), (0.65, 0.28, 0.2))]
              + [("csphere", (sx * 0.3, sy * 0.12, sz * 0.08 + 0.02), 0.04) for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)],
    )
    legs = [("LF", 1.0, 1.0), ("LH", -1.0, 1.0), ("RF", 1.0, -1.0), ("RH", -1.0, -1.0)]
    jd = dict(stiffness=0.0, damping=0.0, armature=0.0)
    haa, free = math.degrees(0.72), math.degrees(9.42)
    hips, thighs, shanks = [], [], []
    for nm, sx, sy in legs:
        hips.append(dict(
            name=f"{nm}_HIP", parent="base", pos=(sx * 0.2999, sy * 0.104, 0.0), quat=(1, 0, 0, 0), mass=2.781,
            joints=[dict(name=f"{nm}_HAA", anchor=(0, 0, 0), axis=(1, 0, 0), range=(-haa, haa), **jd)],
            geoms=[("sphere", (sx * 0.0599, sy * 0.04, 0.0), 0.08)],
        ))
        thighs.append(dict(
            name=f"{nm}_THIGH", parent=f"{nm}_HIP", pos=(sx * 0.0599, sy * 0.08381, 0.0), quat=(1, 0, 0, 0), mass=3.071,
            joints=[dict(name=f"{nm}_HFE", anchor=(0, 0, 0), axis=(0, 1, 0), range=(-free, free), **jd)],
            geoms=[("capsule", (0, sy * 0.05, 0), (0, sy * 0.09, -0.285), 0.04)],
        ))
        shanks.append(dict(
            name=f"{nm}_SHANK", parent=f"{nm}_THIGH", pos=(0.0, sy * 0.1003, -0.285), quat=(1, 0, 0, 0), mass=0.838,
            joints=[dict(name=f"{nm}_KFE", anchor=(0, 0, 0), axis=(0, 1, 0), range=(-free, free), **jd)],
            geoms=[("capsule", (0, 0, 0), (sx * 0.08795, sy * 0.01305, -0.33797), 0.02),
                   ("sphere", (sx * 0.08795, sy * 0.01305, -0.33797), 0.03)],
        ))
    # document order = breadth-first: the four hips, then thighs, then shanks (the reference's DoF order)
    return [base] + hips + thighs + shanks


ANYMAL = dict(
    name="Anymal",
    bodies=anymal_bodies,
    density=1000.0,                       # overridden per body by the explicit masses
    spawn_height=0.62,                    # tasks/anymal.py:100
    sensor_bodies=[],
    dof_order_in_body={},
    gears={"HAA": 0.0, "HFE": 0.0, "KFE": 0.0},  # no effort actuation: PD position targets only (anymal.py:190)
    # set_drive(..., "angular", "position", 0, 400, 40, 1000) (tasks/anymal.py:128, tasks/utils/usd_utils.py:71-82).
    # UsdPhysics angular drives are specified PER DEGREE (stiffness in N*m/deg, damping in N*m*s/deg); the joint
    # state and targets of the tensor API are in radians, hence the 180/pi factors.  The max force is 1000 N*m.
    drive=dict(stiffness=400.0 * 57.29577951308232, damping=40.0 * 57.29577951308232, max_force=1000.0),
)

ROBOTS = {"Humanoid": HUMANOID, "Ant": ANT, "Cartpole": CARTPOLE, "Anymal": ANYMAL}
