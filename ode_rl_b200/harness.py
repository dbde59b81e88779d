"""The torch-side CALLER of the hot path, for bench.py, smoke() and the parity tests: the same wiring as the
reference's models/ODEConvGRU.py:12-99 (ConvEncode -> ODEConvGRUCell -> DiffEqSolver -> ConvDecode -> sigmoid)
with identical submodule / parameter names, so a reference state_dict loads with strict=True.  ConvEncode /
ConvDecode stay plain torch (north_star); only the two starred modules are ode_rl_b200's.  Where the reference
tree is available, models/ODEConvGRU.py itself can be used instead: this class exists because
/root/reference does not travel to the GPU box.
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from .modules import bn_relu, bn_relu_supported, warp_compose, warp_compose_supported, encode_stem_supported, encode_stem_then_conv, conv_same_padded, conv_same_padded_supported, DiffEqSolver, ODEConvGRUCell, ODEFunc, conv3x3s2, conv3x3s2_supported, conv_transpose4x4s2_phases, conv_transpose4x4s2_supported, conv_same, conv_same_supported, decode_tail, decode_tail_supported, phases_to_nchw, upsample2x, upsample2x_supported


class _Encoder(nn.Module):          # models/ODEConvGRU.py:101-119 with n_downs == 2
    def __init__(self, n_inputs, out_ch):
        super().__init__()
        act = nn.LeakyReLU(negative_slope=0.2, inplace=True)
        self.encoder = nn.Sequential(nn.Conv2d(n_inputs, 16, 3, 2, 1), act, nn.Conv2d(16, out_ch, 3, 2, 1), act)
        self.fast_second_layer = os.environ.get("ODEX_TORCH_ENCODER") is None

    def forward(self, x):
        e = self.encoder
        if self.fast_second_layer:
            if encode_stem_supported(x, e[0], e[1], e[2]):
                # the 1 -> 16 stem + LeakyReLU as one HBM-bound kernel that writes the space-to-depth cells the strided
                # tensor-core convolution reads (no cuDNN SGEMM / activation / shuffle passes; backward: dw, db in one kernel)
                return nn.functional.leaky_relu(encode_stem_then_conv(x, e[0], e[1], e[2]), e[3].negative_slope)
            h = e[1](e[0](x))
            if conv3x3s2_supported(h, e[2]):
                # the 16 -> 64 strided convolution as a 3x3 convolution over space-to-depth cells on the
                # tcgen05 path; its NHWC result is what the encoder recurrence reads
                return nn.functional.leaky_relu(conv3x3s2(h, e[2]), e[3].negative_slope)     # not in place: the result is a view
            return e[3](e[2](h))
        return self.encoder(x)


class _Decoder(nn.Module):          # models/ODEConvGRU.py:122-140 with n_ups == 2
    def __init__(self, n_inputs, out_ch):
        super().__init__()
        act = nn.LeakyReLU(negative_slope=0.2, inplace=True)
        self.decoder = nn.Sequential(nn.ConvTranspose2d(n_inputs, 32, 4, 2, 1), act, nn.ConvTranspose2d(32, out_ch, 4, 2, 1))
        self.fast_first_layer = os.environ.get("ODEX_TORCH_DECODER") is None

    def forward(self, x):
        d0 = self.decoder[0]
        if self.fast_first_layer and conv_transpose4x4s2_supported(x, d0):
            # the 64 -> 32 transposed convolution (60 % of the torch-side time under exact-fp32 cuDNN) as a
            # 3x3 convolution to 4 x 32 phase channels on the tcgen05 path; same parameters, fp32-level result
            P = conv_transpose4x4s2_phases(x, d0)                      # phase-blocked, before the activation
            if not P.is_contiguous():
                P = P.contiguous()
            if decode_tail_supported(P, self.decoder[1], self.decoder[2]):
                return decode_tail(P, self.decoder[1], self.decoder[2])    # LeakyReLU + ConvTranspose2d(32, 1, 4, 2, 1), one kernel
            return self.decoder[2](self.decoder[1](phases_to_nchw(P)))
        return self.decoder(x)


class ODEConvGRUHarness(nn.Module):
    def __init__(self, method="euler", rtol=1e-4, atol=1e-5, in_channels=1, resolution=64, ch=64, n_units=64,
                 n_ode_layers=3, device=torch.device("cpu")):
        super().__init__()
        self.device = device
        hw = (resolution // 4, resolution // 4)
        self.conv_encoder = _Encoder(in_channels, ch).to(device)
        self.ode_encoder_func = ODEFunc(n_inputs=ch, n_outputs=ch, n_layers=n_ode_layers, n_units=n_units,
                                        downsize=False, nonlinear='relu', device=device, final_act=False)
        self.ode_convgru_cell = ODEConvGRUCell(self.ode_encoder_func, None, hw, ch, device=device)
        self.ode_decoder_func = ODEFunc(n_inputs=ch, n_outputs=ch, n_layers=n_ode_layers, n_units=n_units,
                                        downsize=False, nonlinear='relu', device=device, final_act=False)
        self.diffeq_solver = DiffEqSolver(self.ode_decoder_func, method, odeint_rtol=rtol, odeint_atol=atol, device=device)
        self.conv_decoder = _Decoder(ch, in_channels).to(device)

    def forward(self, inputs, observed_tp, tp_to_predict, return_latents=True):
        b, t, c, h, w = inputs.size()
        # models/ODEConvGRU.py:63-66 encodes the frames in batch-major order and permutes the FEATURES to time-first; the encoder
        # is per frame, so the same result comes from permuting the (64x smaller) frames instead: the recurrence then reads its
        # [T, B, ...] input in place and its input gradient goes back without a copy (2 x 84 MB per step at B = 128)
        enc = self.conv_encoder(inputs.transpose(0, 1).reshape(t * b, c, h, w))
        _, c_, h_, w_ = enc.size()
        enc = enc.view(t, b, c_, h_, w_)                                     # time first
        mean_z0, std_z0 = self.ode_convgru_cell(enc, observed_tp)
        sol_y = self.diffeq_solver(mean_z0, tp_to_predict)                  # [T,B,C,H,W]
        T, B, C, H, W = sol_y.size()
        pred = torch.sigmoid(self.conv_decoder(sol_y.reshape(B * T, C, H, W)))
        pred = pred.view(T, B, *pred.shape[1:]).permute(1, 0, 2, 3, 4)
        if return_latents:
            return {"pred": pred, "sol": sol_y, "mean_z0": mean_z0, "std_z0": std_z0}
        return pred


class _VidEncoder(nn.Module):        # models/VidODE.py:11-26 (BatchNorm conv encoder; stays torch)
    def __init__(self, input_dim=3, ch=64, n_downs=2, device=None):
        super().__init__()
        layers = [nn.Conv2d(input_dim, ch, 3, 1, 1), nn.BatchNorm2d(ch), nn.ReLU()]
        for _ in range(n_downs):
            layers += [nn.Conv2d(ch, ch * 2, 4, 2, 1), nn.BatchNorm2d(ch * 2), nn.ReLU()]
            ch *= 2
        self.cnn_encoder = nn.Sequential(*layers).to(device)
        self.fast_convs = os.environ.get("ODEX_TORCH_ENCODER") is None

    def forward(self, x):
        if not (self.fast_convs and x.is_cuda):
            return self.cnn_encoder(x)
        # the 4x4 stride-2 down-sampling convolutions (32 -> 64 -> 128 channels on B*T images) as 3x3 convolutions over
        # space-to-depth cells on the tensor-core path (modules.conv3x3s2); the 3-channel stem, BatchNorm and ReLU stay torch
        # channels-last frames: cuDNN's NHWC stem + BatchNorm kernels (the NCHW BatchNorm kernel alone was 3.2 ms of the
        # forward at 320 x 32 x 128 x 128; measured 38.9 -> 36.0 ms per forward)
        x = x.contiguous(memory_format=torch.channels_last)
        mods = list(self.cnn_encoder)
        i = 0
        while i < len(mods):
            m = mods[i]
            if i + 1 < len(mods) and isinstance(mods[i + 1], nn.ReLU) and bn_relu_supported(x, m):
                x = bn_relu(x, m)               # BatchNorm (batch statistics) + ReLU: one statistics pass + one apply pass
                i += 2
                continue
            x = conv3x3s2(x, m) if conv3x3s2_supported(x, m) else m(x)
            i += 1
        return x


class _VidDecoder(nn.Module):        # models/VidODE.py:28-44 (bilinear 2x upsample + conv/BN/ReLU; stays torch)
    def __init__(self, input_dim=256, output_dim=3, n_ups=2, device=None):
        super().__init__()
        layers, ch = [], input_dim
        for _ in range(n_ups):
            layers += [nn.Upsample(scale_factor=2, mode='bilinear', align_corners=False),
                       nn.Conv2d(ch, ch // 2, 3, 1, 1), nn.BatchNorm2d(ch // 2), nn.ReLU()]
            ch //= 2
        layers += [nn.Conv2d(ch, output_dim, 3, 1, 1)]
        self.cnn_decoder = nn.Sequential(*layers).to(device)
        self.fast_convs = os.environ.get("ODEX_TORCH_DECODER") is None

    def forward(self, x):
        if not (self.fast_convs and x.is_cuda):
            return self.cnn_decoder(x)
        # the wide 3x3 convolutions (256 -> 128 on 64x64, 128 -> 64 on 128x128) on the tensor-core
        # convolution of the path; upsample / BatchNorm / ReLU and the 6-channel head stay torch, in channels-last memory
        x = x.contiguous(memory_format=torch.channels_last)
        mods = list(self.cnn_decoder)
        i = 0
        while i < len(mods):
            m = mods[i]
            if i + 1 < len(mods) and isinstance(mods[i + 1], nn.ReLU) and bn_relu_supported(x, m):
                x = bn_relu(x, m)
                i += 2
                continue
            if upsample2x_supported(x, m):
                x = upsample2x(x)
            elif conv_same_supported(x, m):
                x = conv_same(x, m)
            elif conv_same_padded_supported(x, m):
                x = conv_same_padded(x, m)          # the 64 -> in_channels + 3 head, output channels padded to 64
            else:
                x = m(x)
            i += 1
        return x


class VidODEHarness(nn.Module):
    """BASELINE configs[4]: the wiring of the reference's models/VidODE.py:46-141 (BN conv encoder -> ODEConvGRUCell at
    32 * 2**n_downs = 128 channels -> DiffEqSolver -> per-step conv decoder on cat(z_t, z_{t-1}) -> recursive flow warp ->
    mask compose) with the same constructor order (the same seed gives the same initial weights; checked against the
    reference class in tests/test_golden.py) and the same submodule / parameter names.  Only the two starred modules
    are ode_rl_b200's; encoder, decoder, grid_sample and the composition stay torch.
    Two harness fixes, not path changes: the reference hands the time-first cell a batch-first
    tensor (models/VidODE.py:101-105) and views the time-first solution as batch-first (:109-110); the harness permutes
    in both places, as upstream Vid-ODE does.  oracle/make_golden.py applies the same two permutes around the
    reference's own submodules to produce tests/golden/vidode_*.pt."""

    def __init__(self, method="dopri5", in_channels=3, resolution=128, n_downs=2, n_layers=2, rtol=1e-4, atol=1e-5,
                 device=torch.device("cpu")):
        super().__init__()
        self.device, self.in_channels = device, in_channels
        self.resize = 2 ** n_downs
        hw = (resolution // self.resize, resolution // self.resize)
        ch = 32
        in_dim = out_dim = ch * self.resize
        self.conv_encoder = _VidEncoder(input_dim=in_channels, ch=ch, n_downs=n_downs, device=device)
        self.ode_encoder_func = ODEFunc(n_inputs=in_dim, n_outputs=out_dim, n_layers=n_layers, n_units=in_dim // 2,
                                        downsize=False, nonlinear='relu', final_act=False, device=device)
        self.encoder_z0 = ODEConvGRUCell(self.ode_encoder_func, None, hw, out_dim, device=device)
        self.ode_decoder_func = ODEFunc(n_inputs=in_dim, n_outputs=out_dim, n_layers=n_layers, n_units=in_dim // 2,
                                        downsize=False, nonlinear='relu', final_act=False, device=device)
        self.diffeq_solver = DiffEqSolver(self.ode_decoder_func, method, odeint_rtol=rtol, odeint_atol=atol, device=device)
        self.conv_decoder = _VidDecoder(out_dim * 2, in_channels + 3, n_downs, device=device)
        self.fast_warp = os.environ.get("ODEX_TORCH_WARP") is None

    def latents(self, inputs, batch_dict):
        """The hot path alone: frames -> (mean_z0, sol_y [B,T,C,H',W'])."""
        b, t, c, h, w = inputs.size()
        # frames permuted to time-major before the (per-frame, batch-statistics) encoder instead of its features after it
        enc = self.conv_encoder(inputs.transpose(0, 1).reshape(t * b, c, h, w))
        enc = enc.view(t, b, *enc.shape[-3:])                                 # time first
        mean_z0, _ = self.encoder_z0(enc, batch_dict['observed_tp'], batch_dict.get("observed_mask"))
        sol_y = self.diffeq_solver(mean_z0, batch_dict['tp_to_predict'])      # [T,B,C,H',W']
        return mean_z0, sol_y.permute(1, 0, 2, 3, 4)

    def forward(self, inputs, batch_dict):
        b, t, c, h, w = inputs.size()
        mean_z0, sol_y = self.latents(inputs, batch_dict)
        n_pred = sol_y.size(1)
        # models/VidODE.py:143-160: decoder on cat(z_t, z_{t-1}); z_{-1} = the embedding of the last observed frame
        prev = self.conv_encoder(inputs[:, -1, ...])
        outs = []
        for k in range(n_pred):
            cur = sol_y[:, k, ...]
            outs.append(self.conv_decoder(torch.cat([cur, prev], dim=1)))
            prev = cur
        heads = outs
        outs = torch.stack(outs, dim=1)                                       # [B,T,2+c+1(+...),h,w]
        flows, inter = outs[:, :, :2, ...], outs[:, :, 2:2 + c, ...]
        masks = torch.sigmoid(outs[:, :, 2 + c:, ...])
        # models/VidODE.py:121-128,162-186: warp the last observed frame recursively along the predicted flow
        gx = torch.linspace(-1.0, 1.0, w, device=inputs.device).view(1, 1, w, 1).expand(b, h, -1, -1)
        gy = torch.linspace(-1.0, 1.0, h, device=inputs.device).view(1, h, 1, 1).expand(b, -1, w, -1)
        grid = torch.cat([gx, gy], 3).float()
        frame, warped = inputs[:, -1, ...], []
        if self.fast_warp and warp_compose_supported(frame, heads[0]):
            # the recursive warp + the composition, one kernel per predicted frame, read straight from the head's NHWC output
            lin_x, lin_y = torch.linspace(-1.0, 1.0, w, device=inputs.device), torch.linspace(-1.0, 1.0, h, device=inputs.device)
            preds = []
            for k in range(n_pred):
                frame, pk = warp_compose(frame, heads[k], lin_x, lin_y)
                warped.append(frame)
                preds.append(pk)
            warped = torch.stack(warped, dim=1)
            pred_x = torch.stack(preds, dim=1).view(b, -1, c, h, w)
            extra = {"optical_flow": flows, "warped_pred_x": warped, "pred_intermediates": inter, "pred_masks": masks,
                     "mean_z0": mean_z0, "sol_y": sol_y}
            return pred_x, extra
        for k in range(n_pred):
            fl = torch.cat([flows[:, k, 0:1] / ((w - 1.0) / 2.0), flows[:, k, 1:2] / ((h - 1.0) / 2.0)], dim=1)
            frame = nn.functional.grid_sample(frame, grid + fl.permute(0, 2, 3, 1), padding_mode="border", align_corners=False)   # the default the reference runs with
            warped.append(frame)
        warped = torch.stack(warped, dim=1)
        pred_x = (masks * warped + (1 - masks) * inter).view(b, -1, c, h, w)
        extra = {"optical_flow": flows, "warped_pred_x": warped, "pred_intermediates": inter, "pred_masks": masks,
                 "mean_z0": mean_z0, "sol_y": sol_y}
        return pred_x, extra


class GraphedTrainStep:
    """One training step (zero_grad -> forward -> loss -> backward -> [gradient all-reduce] -> optimizer.step) captured once as
    a CUDA graph and replayed: the step is ~700 kernel launches enqueued from Python through ctypes, i.e. ~17 ms of host time
    per step, which a graph replay removes.  Fixed-step solvers only (the adaptive solver reads its controller state back on
    the host once per forward, which cannot be captured).
    `grad_hook` (the NCCL flat-bucket all-reduce of the multi-GPU run) is captured INSIDE the graph together with the optimizer
    step, so that nothing between the last weight-gradient kernel and Adam passes through the host (round 1 replayed the graph,
    then ran the all-reduce and the optimizer eagerly: a graph split and ~0.24 ms per step at 8 GPUs).  If the collective cannot
    be captured on this build (`ODEX_GRAPH_NCCL=0`, or the capture raises), it falls back to that eager tail."""

    def __init__(self, model, optimizer, loss_fn, x, y, obs, tp, grad_hook=None, warmup=3):
        self.model, self.opt, self.loss_fn, self.grad_hook = model, optimizer, loss_fn, grad_hook
        self.x, self.y = x.clone(), y.clone()                 # static input buffers
        self.obs, self.tp = obs, tp
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):                            # allocator / cuDNN autotune / weight-pack / NCCL warm-up
                self._eager()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        from . import _lib
        self.hook_in_graph = grad_hook is not None and os.environ.get("ODEX_GRAPH_NCCL", "1") != "0"
        for attempt in range(2):
            n0 = _lib.launch_count()
            self.graph = torch.cuda.CUDAGraph()
            self.opt.zero_grad(set_to_none=True)
            try:
                with torch.cuda.graph(self.graph):
                    self.loss = self._fwd_bwd()
                    if self.grad_hook is None:
                        self.opt.step()
                    elif self.hook_in_graph:
                        self.grad_hook()
                        self.opt.step()
                break
            except Exception:                                  # a collective that cannot be captured: eager tail instead
                if not self.hook_in_graph or attempt == 1:
                    raise
                self.hook_in_graph = False
                torch.cuda.synchronize()
        self.launches_per_step = _lib.launch_count() - n0      # libodex kernels inside the graph

    def _fwd_bwd(self):
        self.opt.zero_grad(set_to_none=True)
        pred = self.model(self.x, self.obs, self.tp, return_latents=False)
        loss = self.loss_fn(pred, self.y)
        loss.backward()
        return loss

    def _eager(self):
        loss = self._fwd_bwd()
        if self.grad_hook is not None:
            self.grad_hook()
        self.opt.step()
        return loss

    def __call__(self, x=None, y=None):
        if x is not None:
            self.x.copy_(x, non_blocking=True)
        if y is not None:
            self.y.copy_(y, non_blocking=True)
        self.graph.replay()
        if self.grad_hook is not None and not self.hook_in_graph:
            self.grad_hook()
            self.opt.step()
        return self.loss.detach()
