"""Latent interface either side of the INN: what the reference's LightningModules do
between their frozen encoders / decoders and ``flow(x, cond, reverse=...)``.

Mirrors, with the same argument meaning, RNG consumption order and tensor shapes,
  * ``PokeMotionModelFC.make_flow_input / forward_sample / forward_density``
    (models/second_stage_video_fc.py:273-345, 356-378, 380-387) -> :class:`SecondStageLatentInterface`
  * ``FCAEINNModel.forward_sample / forward_density`` (experiments/fully_connected_ae_INN.py:104-127)
    -> :class:`FlowAELatentInterface`
  * the third stage, poke-conditional and unconditional: ``make_flow_input / forward_sample / forward_density`` and the
    frozen second stage used as feature extractor (models/third_stage_video_fc_conditional.py:180-255,
    models/third_stage_video_fc.py:166-231) -> :class:`ThirdStageLatentInterface`

The encoders / decoders themselves (CNN poke and image encoders, GRU video decoder, BigAE) are outside the hot path
(SURVEY.md section 8f, N2): they are passed in as callables.  What changes against the reference is the sampling loop:
the reference re-encodes the same image + poke and calls the flow once per draw; here the conditioning is encoded once
and all ``n_samples`` draws go through ONE ``flow.reverse_draws`` call, in which the cond-only subnetworks of the 'cond'
couplings run once per conditioning row (include/ipoke_inn.h, ipoke_inn_reverse_draws).  The latents are drawn exactly
as the reference draws them (one ``torch.randn(shape)`` on the CPU default generator per draw, in draw order), so with
deterministic encoders the samples are those of the reference's loop under the same seed.
"""
import torch


class SecondStageLatentInterface:
    """Host logic of PokeMotionModelFC around its flow (models/second_stage_video_fc.py:273-387).

    flow                  SupervisedTransformer (``flow.flow.reshape`` is read as in :309)
    poke_encoder(poke)    -> poke embedding (``self.poke_embedder.encoder``, :293)
    cond_encoder(X0)      -> image embedding or None (``self.conditioner.encoder``, :300-304); with
                             ``cond_deterministic=False`` it returns a tuple whose 2nd entry is used (:304)
    encode_first_stage(X) -> (motion, mu)  (:389-396)        decode_first_stage(motion, X) -> video (:398-443)
    """

    def __init__(self, flow, poke_encoder, cond_encoder=None, encode_first_stage=None, decode_first_stage=None, *,
                 flow_in_channels, min_spatial_size=1, poke_key="poke", embed_poke_and_image=False,
                 cond_deterministic=True, base_distribution="normal", augment_channels=0, scale_augment=None,
                 shift_augment=None):
        self.flow = flow
        self.poke_encoder = poke_encoder
        self.cond_encoder = cond_encoder
        self.use_cond = cond_encoder is not None
        self.encode_first_stage = encode_first_stage
        self.decode_first_stage = decode_first_stage
        self.flow_in_channels = flow_in_channels
        self.min_spatial_size = min_spatial_size
        self.poke_key = poke_key
        self.embed_poke_and_image = embed_poke_and_image
        self.cond_deterministic = cond_deterministic
        self.base_distribution = base_distribution
        self.augment_input = augment_channels > 0
        self.augment_channels = augment_channels
        self.scale_augment = scale_augment
        self.shift_augment = shift_augment

    # ------------------------------------------------------------------ pieces of make_flow_input
    def encode_condition(self, batch, use_kp_poke=False):
        """(image, poke) -> cond [B, E, 1, 1]: second_stage_video_fc.py:274-306 and :338-343."""
        X = batch["images"]
        if use_kp_poke:
            poke, *_ = batch["keypoint_poke"]                                    # :276-278
            poke = poke.to(torch.float)
        else:
            poke = batch[self.poke_key]
            poke = poke[0] if isinstance(poke, list) else poke                   # :280-281
        if self.embed_poke_and_image:
            poke = torch.cat([poke, X[:, 0]], dim=1)                             # :284
        with torch.no_grad():
            poke_emb = self.poke_encoder(poke)                                   # :293
            if self.use_cond:
                cond = self.cond_encoder(X[:, 0])                                # :301
                if not self.cond_deterministic:
                    _, cond, _ = cond                                            # :304
        cond = torch.cat([cond, poke_emb], dim=1) if self.use_cond else poke_emb  # :338-341
        if cond.dim() != 4:
            cond = cond.unsqueeze(-1).unsqueeze(-1)                              # :342-343
        return cond

    def latent_shape(self, batch_size):
        if self.flow.flow.reshape != "none":                                     # :309-314
            down = self.flow.flow.reshape == "down"
            cn, sp = (4, 2) if down else (1 / 4, .5)
            return [batch_size, int(cn * self.flow_in_channels), int(self.min_spatial_size / sp),
                    int(self.min_spatial_size / sp)]
        return (batch_size, self.flow_in_channels, self.min_spatial_size, self.min_spatial_size)   # :320-321

    def draw_latent(self, X):
        """One reverse-pass input, drawn as the reference draws it (:316, :322, :325-328)."""
        shape = self.latent_shape(X.size(0))
        flow_input = torch.randn(shape).type_as(X).detach()
        if self.base_distribution == "radial":
            flow_input = torch.nn.functional.normalize(flow_input.view(shape[0], -1))
            flow_input = flow_input.T * torch.abs(torch.randn(shape[0]).type_as(X))
            flow_input = flow_input.T.view(shape).detach()
        return flow_input

    # ------------------------------------------------------------------ the reference's three entry points
    def make_flow_input(self, batch, reverse=False, use_kp_poke=False):
        """-> (flow_input, cond), second_stage_video_fc.py:273-345.  reverse=True: a fresh N(0, I) (or radial) latent;
        reverse=False: the first-stage motion encoding (+ learnable-affine noise channels when augment_input, :332-336)."""
        X = batch["images"]
        cond = self.encode_condition(batch, use_kp_poke)
        if reverse:
            flow_input = self.draw_latent(X)
        else:
            with torch.no_grad():
                flow_input, *_ = self.encode_first_stage(X)                      # :330-331
            if self.augment_input:
                aug = torch.randn((flow_input.size(0), self.augment_channels, *flow_input.shape[-2:])).type_as(X)
                aug = self.scale_augment[None, :, None, None] * aug + self.shift_augment[None, :, None, None]
                flow_input = torch.cat([flow_input, aug], dim=1)                 # :334-336
        return flow_input, cond

    def forward_density(self, batch, return_cond=False):
        """-> out, logdet [, cond]  (second_stage_video_fc.py:380-387)."""
        flow_input, cond = self.make_flow_input(batch)
        out, logdet = self.flow(flow_input.detach(), cond, reverse=False)
        if return_cond:
            return out, logdet, cond
        return out, logdet

    def sample_latents(self, batch, n_samples=1, use_keypoint_pokes=False):
        """The INN part of forward_sample for all draws at once -> motion [n_samples, B, C'] (squeezed, augment
        channels dropped: :366-371).  cond is encoded once; latents are drawn per draw in the reference's order."""
        X = batch["images"]
        with torch.no_grad():
            cond = self.encode_condition(batch, use_keypoint_pokes)
            z = torch.stack([self.draw_latent(X) for _ in range(n_samples)])     # [n, B, C, s, s]
            if self.flow.flow.reshape != "none" or self.min_spatial_size != 1:
                raise NotImplementedError("the fused path covers the fully connected latent ([B, C, 1, 1]); spatial "
                                          "latents belong to the MaCow conv flow (SURVEY.md section 8f, N3)")
            out = self.flow.flow.reverse_draws(z, cond, n_samples)               # [n, B, C, 1, 1]
            out = out.squeeze(-1).squeeze(-1)
            if self.augment_input:
                out = out[..., :-self.augment_channels]
        return out

    def forward_sample(self, batch, n_samples=1, n_logged_vids=1, show_progress=False, add_first_frame=False,
                       use_keypoint_pokes=False):
        """-> list of n_samples decoded videos (first n_logged_vids of the batch, on the CPU), as
        second_stage_video_fc.py:356-378 (show_progress is accepted for signature parity: there is no per-draw loop)."""
        X = batch["images"]
        motion = self.sample_latents(batch, n_samples, use_keypoint_pokes)
        video_samples = []
        with torch.no_grad():
            for n in range(n_samples):
                out_video = self.decode_first_stage(motion[n], X)                # :372
                if add_first_frame:
                    out_video = torch.cat([X[:, 0].unsqueeze(1), out_video], dim=1)   # :373-374
                video_samples.append(out_video[:n_logged_vids].cpu())            # :376
        return video_samples


class FlowAELatentInterface:
    """Host logic of FCAEINNModel around its INN (experiments/fully_connected_ae_INN.py:104-127): ``ae`` is the frozen
    flow autoencoder (``encode(x)`` returns a distribution with ``.sample()`` / ``.mode()``, ``decode(z)`` an image)."""

    def __init__(self, INN, ae):
        self.INN = INN
        self.ae = ae

    def forward_density(self, batch):
        X = batch["flow"]
        with torch.no_grad():
            encv = self.ae.encode(X).sample()                                    # :119-120
        return self.INN(encv, reverse=False)                                     # :125

    def forward_sample(self, batch, n_samples=1, n_logged_imgs=1):
        """The reference encodes the batch once per draw only to get the latent's shape (:108-110); here once."""
        image_samples = []
        with torch.no_grad():
            like = self.ae.encode(batch["flow"]).mode()
            z = torch.stack([torch.randn_like(like) for _ in range(n_samples)])  # :110, draw order kept
            B = like.shape[0]
            out = self.INN(z.reshape(n_samples * B, *like.shape[1:]), reverse=True)   # unconditional: rows independent
            out = out.reshape(n_samples, B, *out.shape[1:])
            for n in range(n_samples):
                image_samples.append(self.ae.decode(out[n])[:n_logged_imgs])     # :112-113
        return image_samples


class ThirdStageLatentInterface:
    """Host logic of the third stage around its INN: the poke-conditional model
    (models/third_stage_video_fc_conditional.py:180-255, INN = SupervisedTransformer, cond = the poke half of the
    frozen second stage's conditioning) and the unconditional one (models/third_stage_video_fc.py:166-231, INN =
    UnsupervisedTransformer3; pass ``cond=None``).

    INN            the flow (called as ``INN(x, cond, reverse=...)`` / ``INN(x, reverse=...)``)
    flow_encoder   frozen optical-flow autoencoder: ``encode(X).sample()``, ``decode(z)``
    second_stage   object with ``forward_density(batch, return_cond=True)`` (a :class:`SecondStageLatentInterface` or
                   the reference's PokeMotionModelFC), used as feature extractor (:241-247)
    flow_in_channels / flow_z_dim   config['architecture']['flow_in_channels'], flow_enc_config[...]['z_dim']
    batch_size     config['data']['batch_size'] (the reverse latent takes ITS row count, not the batch's: :183)
    poke_z_dim     second_stage.poke_embedder.config['architecture']['z_dim'] (:246)
    """

    def __init__(self, INN, flow_encoder, second_stage=None, *, flow_in_channels, flow_z_dim, batch_size, poke_z_dim=0,
                 base_distribution="normal"):
        if flow_in_channels < flow_z_dim:
            raise AssertionError(f"{flow_in_channels} < {flow_z_dim}")           # :193
        self.INN = INN
        self.flow_encoder = flow_encoder
        self.second_stage = second_stage
        self.flow_in_channels = flow_in_channels
        self.flow_z_dim = flow_z_dim
        self.batch_size = batch_size
        self.poke_z_dim = poke_z_dim
        self.base_distribution = base_distribution

    def make_flow_input(self, batch, reverse=False):
        """third_stage_video_fc_conditional.py:180-203.  reverse=True: N(0, I) (or radial) latent of
        [batch_size, flow_in_channels, 1, 1] drawn on X's device; reverse=False: the flow-AE latent padded with fresh
        N(0, I) channels up to flow_in_channels."""
        X = batch["flow"]
        if reverse:
            shape = [self.batch_size, self.flow_in_channels, 1, 1]
            flow_input = torch.randn(shape, dtype=X.dtype, layout=X.layout, device=X.device).detach()   # :184
            if self.base_distribution == "radial":
                flow_input = torch.nn.functional.normalize(flow_input.view(shape[0], -1))               # :186-189
                flow_input = flow_input.T * torch.abs(torch.randn(shape[0]).type_as(X))
                flow_input = flow_input.T.view(shape).detach()
            return flow_input
        with torch.no_grad():
            enc_vec = self.flow_encoder.encode(X).sample()                       # :191
            rest = list(enc_vec.shape)
            rest[1] = self.flow_in_channels - self.flow_z_dim                    # :192-193
            if rest[1] != 0:
                motion_rest = torch.randn(rest, dtype=enc_vec.dtype, layout=enc_vec.layout, device=enc_vec.device)
                return torch.cat((enc_vec, motion_rest), 1)                      # :196-197
            return enc_vec

    def poke_encoding(self, cond):
        """The poke half of the second stage's conditioning [image emb | poke emb] (:246)."""
        return cond[:, -self.poke_z_dim:]

    def forward_density(self, batch, cond=None):
        """-> out, logdet, optical_flow_vector (a detached copy of the INN's input: the INN must not mutate it, :237)."""
        flow_input = self.make_flow_input(batch, reverse=False)
        optical_flow_vector = flow_input.clone().detach()
        if cond is None:
            out, logdet = self.INN(flow_input.detach(), reverse=False)           # third_stage_video_fc.py:224
        else:
            out, logdet = self.INN(flow_input.detach(), cond, reverse=False)     # :238
        return out, logdet, optical_flow_vector

    def density_with_second_stage(self, batch):
        """The model part of training_step (:241-247; unconditional: third_stage_video_fc.py:229-231):
        -> out, logdet, optical_flow_vector, out_hat (the frozen second stage's latent, the reconstruction target)."""
        conditional = self.poke_z_dim > 0
        with torch.no_grad():
            if conditional:
                out_hat, _, cond = self.second_stage.forward_density(batch, return_cond=True)
            else:
                out_hat, _ = self.second_stage.forward_density(batch)
        pe = self.poke_encoding(cond).detach() if conditional else None
        out, logdet, vec = self.forward_density(batch, pe)
        return out, logdet, vec, out_hat

    def forward_sample(self, batch, cond=None, n_samples=1, n_logged_imgs=1, flow_input=None):
        """-> (list of n_samples decoded optical-flow images, residual latent channels of the last draw), as :215-233.
        All draws go through ONE reverse call (the reference calls the INN once per draw): ``reverse_draws`` on the
        shared conditioning for the conditional INN, one stacked batch for the unconditional one.  Latents are drawn
        per draw in the reference's order; a given ``flow_input`` replaces the single draw (:222-223)."""
        if flow_input is not None:
            assert n_samples == 1, "second pass will random sample and overwrite given input"
        opt_flow_dim = self.flow_z_dim
        image_samples = []
        with torch.no_grad():
            zs = [flow_input if (flow_input is not None and n == 0) else self.make_flow_input(batch, reverse=True)
                  for n in range(n_samples)]
            z = torch.stack([t[:n_logged_imgs] for t in zs])                      # [n, L, C, 1, 1] (or [n, L, C])
            n, L = z.shape[0], z.shape[1]
            if cond is None:
                out = self.INN(z.reshape(n * L, *z.shape[2:]), reverse=True)     # rows are independent
                out = out.reshape(n, L, *out.shape[1:])
            elif n == 1:
                out = self.INN(z[0], cond[:n_logged_imgs], reverse=True)[None]
            else:
                out = self.INN.flow.reverse_draws(z, cond[:n_logged_imgs], n)
                out = out.reshape(n, L, -1, 1, 1)
            residual = None
            for k in range(n):
                o, residual = out[k][:, :opt_flow_dim], out[k][:, opt_flow_dim:]   # :226
                if o.dim() == 4:
                    o = o.squeeze(-1).squeeze(-1)
                image_samples.append(self.flow_encoder.decode(o))                # :229
        return image_samples, residual
