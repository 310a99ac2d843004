"""volume_render -- signature and tuple order of the reference's view_synthesis/nerf/volumetric_render.py:36-66."""
from codenerf_b200 import ops


def volume_render(radiance_field, depth_values, ray_directions):
    """radiance_field [..., S, 4], depth_values [..., S], ray_directions [..., 3]
    -> (rgb_map, disp_map, acc_map, weights, depth_map), differentiable in all five."""
    lead = depth_values.shape[:-1]
    S = depth_values.shape[-1]
    rgb, disp, acc, weights, depth = ops.composite(radiance_field.reshape(-1, S, 4), depth_values.reshape(-1, S),
                                                   ray_directions.reshape(-1, 3))
    return (rgb.reshape(*lead, 3), disp.reshape(lead), acc.reshape(lead), weights.reshape(*lead, S),
            depth.reshape(lead))
