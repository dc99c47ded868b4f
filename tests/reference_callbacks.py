"""The reference's example right-hand side callbacks as the GPU tests use them (the GPU box has no
``/root/reference``).  ``tests/test_reference_source_pins.py`` traces these and the functions taken from the
reference's own files and requires the generated device functors to be identical, character for character."""


def van_der_pol(ozone_vars):                     # ozone/paper_examples/van_der_pol/main_model.py:6-14
    x0 = ozone_vars.states['x0']
    x1 = ozone_vars.states['x1']
    u = ozone_vars.dynamic_parameters['u']
    ozone_vars.d_states['x0'] = ((1.0 + (- 1.0) * (x1**2.0))*x0 + (-1*x1) + u).reshape(-1, 1)
    ozone_vars.d_states['x1'] = (x0*1.0).reshape(-1, 1)
    ozone_vars.d_states['J'] = (x0**2.0 + x1**2.0 + u**2.0).reshape(-1, 1)


def lunar_ascent(ozone_vars):                    # ozone/paper_examples/ascent_system/main_model.py:6-21
    import numpy as np
    g0, R0, vex, w2, T = 1.625, 1738100, 3049.87, 1., 1.8335634
    T_m = T/ozone_vars.states['m']
    n = ozone_vars.num_nodes
    ozone_vars.d_states['rx'] = ozone_vars.states['Vx'].reshape(-1, 1)
    ozone_vars.d_states['ry'] = ozone_vars.states['Vy'].reshape(-1, 1)
    ozone_vars.d_states['rz'] = ozone_vars.states['Vz'].reshape(-1, 1)
    ozone_vars.d_states['Vx'] = (-w2*ozone_vars.states['rx'] + (T_m) * ozone_vars.dynamic_parameters['ux']).reshape(-1, 1)
    ozone_vars.d_states['Vy'] = (-w2*ozone_vars.states['ry'] + (T_m) * ozone_vars.dynamic_parameters['uy']).reshape(-1, 1)
    ozone_vars.d_states['Vz'] = (-w2*ozone_vars.states['rz'] + (T_m) * ozone_vars.dynamic_parameters['uz']).reshape(-1, 1)
    ozone_vars.d_states['m'] = (-np.ones(n) * (T / vex) * np.sqrt(g0*R0)).reshape(-1, 1)


def lotka_volterra(ozone_vars, c):               # examples/simple_examples/lotka_volterra.py:51-62
    a = ozone_vars.dynamic_parameters['a']
    x = ozone_vars.states['x']
    y = ozone_vars.states['y']
    ozone_vars.d_states['y'] = a*y - 0.5*y*x
    ozone_vars.d_states['x'] = 2.0*x*y - c*x
    ozone_vars.field_outputs['field_x'] = x
    ozone_vars.profile_outputs['x_plus_y'] = x + y


# declaration order of states / dynamic parameters in the reference scripts
# (van_der_pol/main_model.py:44-52, ascent_system/main_model.py:81-94, lotka_volterra.py:42-46)
DECLARATIONS = {
    'van_der_pol': (dict(x0=(), x1=(), J=()), dict(u=())),
    'lunar_ascent': (dict(rx=(), ry=(), rz=(), Vx=(), Vy=(), Vz=(), m=()), dict(ux=(), uy=(), uz=())),
    'lotka_volterra': (dict(y=(), x=()), dict(a=())),
}
CALLBACKS = {'van_der_pol': van_der_pol, 'lunar_ascent': lunar_ascent, 'lotka_volterra': lotka_volterra}
