"""The per-scan-point ray lattice and the post-processing the reference wraps around its kernels."""
import torch


def ray_lattice(camera_pos, theta_range, phi_range, r_range, num_theta, num_phi, num_r):
    """Rays of one scan point: theta/phi linspace grids (indexing 'ij'), d = (sin t cos p, sin t sin p, cos t),
    all origins = camera_pos, radii = linspace(r_range)   (gaussian_model/cuda_autograd.py:254-273).
    -> origins [R,3], dirs [R,3], t_samples [num_r], theta_grid [num_theta,num_phi]"""
    dev = camera_pos.device
    th = torch.linspace(theta_range[0], theta_range[1], num_theta, device=dev)
    ph = torch.linspace(phi_range[0], phi_range[1], num_phi, device=dev)
    tg, pg = torch.meshgrid(th, ph, indexing="ij")
    st = torch.sin(tg)
    dirs = torch.stack([st * torch.cos(pg), st * torch.sin(pg), torch.cos(tg)], -1).reshape(-1, 3).contiguous()
    origins = camera_pos.reshape(1, 3).expand(dirs.shape[0], 3).contiguous()
    t = torch.linspace(r_range[0], r_range[1], num_r, device=dev)
    return origins, dirs, t, tg


def attenuate_and_integrate(rho_density, t_samples, theta_grid, theta_range, phi_range):
    """rho [R,T] -> result [T,A_theta,A_phi] = rho^T / (r^2 + 1e-8) * sin(theta); histogram [T] = sum * dtheta * dphi with
    dtheta = (max-min)/num_theta  (cuda_autograd.py:302-314; the /num, not /(num-1), is the reference's)."""
    n_th, n_ph = theta_grid.shape
    T = t_samples.shape[0]
    res = rho_density.T.reshape(T, n_th, n_ph)
    res = res / (t_samples.view(-1, 1, 1) ** 2 + 1e-8) * torch.sin(theta_grid).unsqueeze(0)
    dth = (theta_range[1] - theta_range[0]) / n_th
    dph = (phi_range[1] - phi_range[0]) / n_ph
    return res, res.sum(dim=(1, 2)) * dth * dph


def angular_range_from_points(input_points):
    """theta/phi min/max of a [Na,5] (x,y,z,theta,phi) lattice with ONE device->host copy
    (the reference does four .item() syncs, rendering_cuda.py:308-311)."""
    a = input_points[:, 3:5]
    v = torch.stack([a.min(0).values, a.max(0).values]).cpu()
    return (float(v[0, 0]), float(v[1, 0])), (float(v[0, 1]), float(v[1, 1]))
