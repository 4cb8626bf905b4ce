"""Mirror of the reference's ``models/`` namespace for the hot path (rendering, nerf, nerf_cls)."""
