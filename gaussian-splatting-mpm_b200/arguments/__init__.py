"""Drop-in for the reference's arguments/__init__.py: the three parameter groups main.py builds from a config JSON
(main.py:338-355), same fields and defaults, same `steps_per_frame = int(frame_dt / substep_dt)`."""
from argparse import ArgumentParser


class GroupParams:
    pass


class ParamGroup:
    """arguments/__init__.py:7-35: every attribute becomes a command-line option whose default is the JSON value"""

    def __init__(self, parser: ArgumentParser, name: str, json_params=None):
        group = parser.add_argument_group(name)
        for key, value in vars(self).items():
            shorthand = key.startswith("_")
            if shorthand:
                key = key[1:]
            t = type(value)
            if json_params and key in json_params:
                value = json_params[key]
            names = ["--" + key] + (["-" + key[0:1]] if shorthand else [])
            if t == bool:
                group.add_argument(*names, default=value, action="store_true")
            else:
                group.add_argument(*names, default=value, type=t)

    def extract(self, args):
        group = GroupParams()
        for k, v in vars(args).items():
            if k in vars(self) or ("_" + k) in vars(self):
                setattr(group, k, v)
        return group


class ModelParams(ParamGroup):
    def __init__(self, parser, json_params=None):
        self.model_path = ""
        self.loaded_iter = -1
        self.debug = False
        super().__init__(parser, "Loading Parameters", json_params)


class MPMParams(ParamGroup):
    def __init__(self, parser, json_params=None):
        self.view_area = []
        self.sim_area = [[-1.0, -1.0, -1.0], [1.0, 1.0, 1.0]]
        self.mask = []
        self.E = 2e6
        self.nu = 0.4
        self.viscosity = 0.05
        self.material = "jelly"
        self.gravity = [0.0, -9.81, 0.0]
        self.density = 1000.0
        self.n_grid = 50
        self.grid_extent = 2.0
        self.substep_dt = 0.0006
        self.frame_dt = 0.03
        self.rotation_degree = [0.0, 0.0, 0.0]
        self.boundary_conditions = []
        self.fitting = False
        super().__init__(parser, "MPM Parameters", json_params)

    def extract(self, args):
        g = super().extract(args)
        g.steps_per_frame = int(g.frame_dt / g.substep_dt)   # arguments/__init__.py:83
        return g


class RenderParams(ParamGroup):
    def __init__(self, parser, json_params=None):
        self.output_path = ""
        self.white_background = False
        self.view_cam_idx = 10
        self.num_frames = 60
        self.save_pcd = False
        self.save_pcd_interval = 10
        super().__init__(parser, "Render Parameters", json_params)


def from_config(config: dict, argv=()):
    """(model_args, sim_args, render_args) as main.py:346-355 extracts them from a loaded config JSON"""
    parser = ArgumentParser(description="Simulation parameters")
    m, s, r = ModelParams(parser, config.get("model")), MPMParams(parser, config.get("mpm")), RenderParams(parser, config.get("render"))
    args = parser.parse_args(list(argv))
    return m.extract(args), s.extract(args), r.extract(args)
