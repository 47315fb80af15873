"""MJCF of the box-narrowphase test scenes (shared by CPU pins and GPU parity tests)."""

BOX_PILE = """
<mujoco>
  <option timestep="0.002"/>
  <worldbody>
    <geom name="floor" type="plane" size="2 2 .1"/>
    <body name="b1" pos="0 0 0.3"><freejoint/><geom name="b1" type="box" size="0.1 0.12 0.08" mass="2.5"/></body>
    <body name="b2" pos="0.05 0.02 0.6"><freejoint/><geom name="b2" type="box" size="0.08 0.06 0.1" mass="1.5"/></body>
    <body name="b3" pos="-0.05 0.1 0.9"><freejoint/><geom name="b3" type="box" size="0.05 0.09 0.07" mass="1.0"/></body>
    <body name="s1" pos="0.1 -0.1 1.2"><freejoint/><geom name="s1" type="sphere" size="0.06" mass="0.5"/></body>
    <body name="c1" pos="-0.1 -0.1 1.5"><freejoint/><geom name="c1" type="capsule" size="0.04 0.09" mass="0.4"/></body>
    <body name="c2" pos="0.15 0.1 1.8"><freejoint/><geom name="c2" type="capsule" size="0.03 0.12" mass="0.3"/></body>
  </worldbody>
</mujoco>"""

BOX_STACK = """
<mujoco>
  <worldbody>
    <geom name="floor" type="plane" size="1 1 1"/>
    <body name="b1" pos="0 0 0.1"><freejoint/><geom type="box" size="0.1 0.1 0.1" mass="2.5"/></body>
    <body name="b2" pos="0.01 0.02 0.3"><freejoint/><geom type="box" size="0.08 0.06 0.1" mass="1.5"/></body>
    <body name="s" pos="0.0 0.0 0.45"><freejoint/><geom type="sphere" size="0.05" mass="0.5"/></body>
    <body name="c" pos="0.4 0 0.05"><freejoint/><geom type="box" size="0.1 0.1 0.05" mass="1"/></body>
    <body name="cap" pos="0.4 0 0.13" euler="0 90 0"><freejoint/><geom type="capsule" size="0.03 0.06" mass="0.3"/></body>
  </worldbody>
</mujoco>"""
