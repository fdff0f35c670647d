"""Densification schedule around the hot path (densification.py:5-56 of the reference).

The per-camera statistics (max_radii2D / xyz_gradient_accum / denom) are accumulated inside the
iteration by ``clmgs_densification_stats`` (engine.backward_one_camera, TrainArgs.densification_stats);
this module decides WHEN the restructuring step runs and calls the model's fused
densify_and_prune / reset_opacity (model.py, csrc/densify.cu).
"""


def check_update_at_this_iter(iteration, bsz, update_interval, update_residual):
    """utils/general_utils.py:130-142: does [iteration, iteration + bsz) contain a multiple of the interval
    (shifted by the residual)?"""
    lo = iteration % update_interval
    hi = lo + bsz
    if lo <= update_residual < hi:
        return True
    return lo <= update_residual + update_interval < hi


def gsplat_densification(iteration, scene, gaussians, batched_screenspace_pkg=None, noise=None, generator=None):
    """Same call as the reference's (densification.py:5).  ``scene`` needs ``cameras_extent``.
    Returns a dict describing what ran (the reference logs the same facts to its log file)."""
    args = gaussians.args
    did = {}
    if args.disable_auto_densification or iteration > args.densify_until_iter:
        return did
    if iteration > args.densify_from_iter and check_update_at_this_iter(iteration, args.bsz,
                                                                         args.densification_interval, 0):
        size_threshold = 20 if iteration > args.opacity_reset_interval else None
        did["densify_and_prune"] = gaussians.densify_and_prune(args.densify_grad_threshold, args.min_opacity,
                                                               scene.cameras_extent, size_threshold, noise=noise,
                                                               generator=generator)
    if check_update_at_this_iter(iteration, args.bsz, args.opacity_reset_interval, 0):
        gaussians.reset_opacity()
        did["reset_opacity"] = True
    return did
