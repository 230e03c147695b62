"""Drop-in modules behind the reference's registry (funasr/register.py:57-65, last registration
wins) for the batched offline Paraformer path.

Each class keeps the reference class's name, constructor keywords, `forward` signature, return
tuple and state_dict keys, so `AutoModel(model="Paraformer", ...)` builds them
by string key and `load_pretrained_model(strict=True)` loads a reference checkpoint unchanged:

    Paraformer             funasr/models/paraformer/model.py:30-596 (`inference` only: one packed engine call)
    CTC                    funasr/models/ctc/model.py:17-247 (`inference` only, over a ConformerEncoder)
    SenseVoiceSmall        funasr/models/sense_voice/model.py:588-906 (`inference` only)
    SenseVoiceEncoderSmall funasr/models/sense_voice/model.py:444-585
    WavFrontend            funasr/frontends/wav_frontend.py:78-158
    SANMEncoder            funasr/models/sanm/encoder.py:188-430
    ConformerEncoder       funasr/models/conformer/encoder.py:285-607 (aishell recipe options: conv2d input layer,
                           legacy rel-pos attention, macaron FFN, convolution module; BASELINE config 4)
    CifPredictorV2         funasr/models/paraformer/cif_predictor.py:172-253
    ParaformerSANMDecoder  funasr/models/paraformer/decoder.py:224-411

The modules exchange the reference's PADDED tensors ([B, T, D] + lengths) at their boundaries;
inside, rows are packed (fb200_pack_rows), run through the CUDA kernels with batch-1 semantics per
utterance, and unpacked again.  There is no CPU path: a forward on CPU tensors raises.

`install()` registers the classes into `funasr.register.tables` under the original keys.
"""
from typing import Dict, List, Optional, Tuple

import torch
from torch import nn

from . import conformer, engine, ops, params, tables

PRECISION = "fp32"  # module-level default; override per module with `precision=` or set_precision()


def set_precision(name: str) -> None:
    global PRECISION
    if name not in engine.ParaformerEngine.PRECISIONS:
        raise ValueError("precision must be one of %s" % sorted(engine.ParaformerEngine.PRECISIONS))
    PRECISION = name


def _require_cuda(t: torch.Tensor, who: str) -> None:
    if not t.is_cuda:
        raise RuntimeError("funasr_b200.%s: input lives on %s -- this path runs on a CUDA device only (no CPU "
                           "fallback); build the model with device='cuda'" % (who, t.device))


class _ParamTree(nn.Module):
    """Holds parameters under the reference's dotted names (e.g. `encoders.3.self_attn.linear_out.weight`)
    so state_dict()/load_state_dict() round-trip reference checkpoints.  Values start as the torch
    default init of the reference layers' shapes."""

    def _declare(self, shapes: Dict[str, Tuple[int, ...]], prefix: str) -> None:
        for name, shape in shapes.items():
            assert name.startswith(prefix)
            parts = name[len(prefix):].split(".")
            mod = self
            for p in parts[:-1]:
                if p not in mod._modules:
                    mod.add_module(p, nn.Module())
                mod = mod._modules[p]
            t = torch.empty(*shape)
            leaf = parts[-1]
            is_norm = any(s.startswith("norm") or s in ("after_norm", "tp_norm") for s in parts[:-1])
            if is_norm:
                nn.init.ones_(t) if leaf == "weight" else nn.init.zeros_(t)
            elif "embed" in parts and leaf == "weight":
                nn.init.normal_(t)
            else:
                wshape = shapes.get(name[: -len(leaf)] + "weight", shape) if leaf == "bias" else shape
                fan_in = 1
                for s in wshape[1:]:
                    fan_in *= s
                bound = 1.0 / max(float(fan_in), 1.0) ** 0.5
                nn.init.uniform_(t, -bound, bound)
            mod.register_parameter(leaf, nn.Parameter(t))
        self._pack = None
        self._pack_key = None

    def _state_key(self):
        return tuple((p.data_ptr(), p._version, str(p.device)) for p in self.parameters())

    def _packed(self, build):
        """Cache of the kernels' weight pack (bf16 splits etc.), rebuilt when parameters change."""
        key = (self._state_key(), PRECISION if getattr(self, "precision", None) is None else self.precision)
        if self._pack is None or self._pack_key != key:
            self._pack = build()
            self._pack_key = key
        return self._pack

    @property
    def _nparts(self) -> int:
        prec = PRECISION if getattr(self, "precision", None) is None else self.precision
        return engine.ParaformerEngine.PRECISIONS[prec]

    def _sd(self, prefix: str) -> Dict[str, torch.Tensor]:
        return {prefix + k: v.detach() for k, v in self.state_dict().items()}


def _lengths_list(lengths) -> List[int]:
    if isinstance(lengths, torch.Tensor):
        return [int(v) for v in lengths.detach().cpu().tolist()]  # one [B] D2H, like make_pad_mask's tolist()
    return [int(v) for v in lengths]


class WavFrontend(nn.Module):
    """funasr/frontends/wav_frontend.py:78-158, on the GPU for the whole batch at once.

    `forward(input [B, Nmax], input_lengths) -> (feats [B, Tmax, n_mels*lfr_m] f32 on the CUDA
    device, feats_lens [B] int64 on the host)`.  `input` may be a host tensor (it is copied to the
    device here: the path's one H2D) or already on the device.  dither != 0 adds N(0, dither^2)
    noise per sample (Kaldi draws it per frame element; parity is defined at dither = 0)."""

    def __init__(self, cmvn_file: str = None, fs: int = 16000, window: str = "hamming", n_mels: int = 80,
                 frame_length: int = 25, frame_shift: int = 10, filter_length_min: int = -1,
                 filter_length_max: int = -1, lfr_m: int = 1, lfr_n: int = 1, dither: float = 1.0,
                 snip_edges: bool = True, upsacle_samples: bool = True, device=None, **kwargs):
        super().__init__()
        if not snip_edges:
            raise ValueError("funasr_b200.WavFrontend: snip_edges=False is not supported")
        self.fs, self.window, self.n_mels = fs, window, n_mels
        self.frame_length, self.frame_shift = frame_length, frame_shift
        self.filter_length_min, self.filter_length_max = filter_length_min, filter_length_max
        self.lfr_m, self.lfr_n = lfr_m, lfr_n
        self.cmvn_file, self.dither, self.snip_edges, self.upsacle_samples = cmvn_file, dither, snip_edges, upsacle_samples
        self.cmvn = None if cmvn_file is None else tables.load_cmvn(cmvn_file)
        self.device = device
        self._pack = None

    def output_size(self) -> int:
        return self.n_mels * self.lfr_m

    def _frontend(self, dev) -> engine.FrontendPack:
        if self._pack is None or self._pack.device != dev:
            self._pack = engine.FrontendPack(dev, cmvn=self.cmvn, window=self.window, n_mels=self.n_mels,
                                             lfr_m=self.lfr_m, lfr_n=self.lfr_n, fs=self.fs,
                                             frame_length=self.frame_length, frame_shift=self.frame_shift,
                                             upscale=self.upsacle_samples)
        return self._pack

    def forward(self, input: torch.Tensor, input_lengths, **kwargs) -> Tuple[torch.Tensor, torch.Tensor]:
        if not torch.cuda.is_available():
            raise RuntimeError("funasr_b200.WavFrontend: no CUDA device (this path has no CPU fallback)")
        dev = input.device if input.is_cuda else torch.device(self.device or ("cuda:%d" % torch.cuda.current_device()))
        lens = _lengths_list(input_lengths)
        fe = self._frontend(dev)
        with torch.cuda.device(dev):
            wave = torch.empty(sum(lens), dtype=torch.float32, device=dev)
            o = 0
            for b, n in enumerate(lens):
                wave[o: o + n].copy_(input[b, :n], non_blocking=True)
                o += n
            if self.dither != 0.0:
                wave.add_(torch.randn_like(wave), alpha=float(self.dither) / (fe.upscale if fe.upscale > 1 else 1.0))
            feats, cu_rows, rows = fe.run(wave, lens)
            out = ops.unpack_rows(feats, cu_rows, max(rows))
        return out, torch.as_tensor(rows, dtype=torch.int64)


class SANMEncoder(_ParamTree):
    """funasr/models/sanm/encoder.py:188-430 (`input_layer="pe"`, SAN-M self-attention, pre-norm)."""

    def __init__(self, input_size: int, output_size: int = 256, attention_heads: int = 4, linear_units: int = 2048,
                 num_blocks: int = 6, dropout_rate: float = 0.1, positional_dropout_rate: float = 0.1,
                 attention_dropout_rate: float = 0.0, input_layer: Optional[str] = "conv2d", pos_enc_class=None,
                 normalize_before: bool = True, concat_after: bool = False, positionwise_layer_type: str = "linear",
                 positionwise_conv_kernel_size: int = 1, padding_idx: int = -1, interctc_layer_idx: List[int] = [],
                 interctc_use_conditioning: bool = False, kernel_size: int = 11, sanm_shfit: int = 0,
                 lora_list: List[str] = None, lora_rank: int = 8, lora_alpha: int = 16, lora_dropout: float = 0.1,
                 selfattention_layer_type: str = "sanm", tf2torch_tensor_name_prefix_torch: str = "encoder",
                 tf2torch_tensor_name_prefix_tf: str = "seq2seq/encoder", precision: Optional[str] = None, **kwargs):
        super().__init__()
        unsupported = []
        if input_layer != "pe":
            unsupported.append("input_layer=%r (only 'pe')" % (input_layer,))
        if selfattention_layer_type != "sanm" or positionwise_layer_type != "linear":
            unsupported.append("attention/positionwise layer types other than sanm/linear")
        if not normalize_before or concat_after or interctc_layer_idx or lora_list:
            unsupported.append("post-norm / concat_after / interctc / lora")
        if sanm_shfit != 0:
            unsupported.append("sanm_shfit != 0")
        if output_size // attention_heads != 128:
            unsupported.append("head dim %d (128 only)" % (output_size // attention_heads))
        if unsupported:
            raise NotImplementedError("funasr_b200.SANMEncoder does not support: " + "; ".join(unsupported))
        self._output_size, self.heads, self.precision = output_size, attention_heads, precision
        self.interctc_layer_idx, self.interctc_use_conditioning, self.conditioning_layer = [], False, None
        self._declare(params.sanm_encoder_shapes(input_size, output_size, linear_units, num_blocks, kernel_size, ""), "")

    def output_size(self) -> int:
        return self._output_size

    def _enc(self, dev) -> engine.EncoderPack:
        return self._packed(lambda: engine.EncoderPack(self._sd("encoder."), dev, heads=self.heads, eps=1e-12,
                                                       nparts=self._nparts))

    def forward(self, xs_pad: torch.Tensor, ilens: torch.Tensor, prev_states: torch.Tensor = None, ctc=None):
        _require_cuda(xs_pad, "SANMEncoder")
        lens = _lengths_list(ilens)
        dev = xs_pad.device
        with torch.cuda.device(dev):
            cu_rows = tables.cu_from_lengths(lens, torch.int32, dev)
            enc = self._enc(dev)
            x = ops.pack_rows(xs_pad.float(), cu_rows, sum(lens))
            x = enc.prologue(x, cu_rows)
            y = enc.run(x, cu_rows, max(lens))
            out = ops.unpack_rows(y, cu_rows, xs_pad.shape[1])
        olens = torch.as_tensor(lens, dtype=torch.int64, device=dev)  # masks.squeeze(1).sum(1)
        return out, olens, None


class SenseVoiceEncoderSmall(_ParamTree):
    """funasr/models/sense_voice/model.py:444-585: the SAN-M stack of SANMEncoder (LayerNorm eps 1e-5) followed by
    `tp_blocks` more blocks and `tp_norm`; `forward(xs_pad, ilens) -> (xs_pad, olens)`.  The caller
    (SenseVoiceSmall.inference, :852-876) has already prepended the 4 query frames."""

    def __init__(self, input_size: int, output_size: int = 256, attention_heads: int = 4, linear_units: int = 2048,
                 num_blocks: int = 6, tp_blocks: int = 0, dropout_rate: float = 0.1, positional_dropout_rate: float = 0.1,
                 attention_dropout_rate: float = 0.0, stochastic_depth_rate: float = 0.0,
                 input_layer: Optional[str] = "conv2d", pos_enc_class=None, normalize_before: bool = True,
                 concat_after: bool = False, positionwise_layer_type: str = "linear",
                 positionwise_conv_kernel_size: int = 1, padding_idx: int = -1, kernel_size: int = 11, sanm_shfit: int = 0,
                 selfattention_layer_type: str = "sanm", precision: Optional[str] = None, **kwargs):
        super().__init__()
        unsupported = []
        if selfattention_layer_type != "sanm" or positionwise_layer_type != "linear":
            unsupported.append("attention/positionwise layer types other than sanm/linear")
        if not normalize_before or concat_after:
            unsupported.append("post-norm / concat_after")
        if sanm_shfit != 0:
            unsupported.append("sanm_shfit != 0")
        if output_size // attention_heads != 128:
            unsupported.append("head dim %d (128 only)" % (output_size // attention_heads))
        if unsupported:
            raise NotImplementedError("funasr_b200.SenseVoiceEncoderSmall does not support: " + "; ".join(unsupported))
        self._output_size, self.heads, self.precision = output_size, attention_heads, precision
        self._declare(params.sanm_encoder_shapes(input_size, output_size, linear_units, num_blocks, kernel_size, "",
                                                 tp_blocks=tp_blocks), "")

    def output_size(self) -> int:
        return self._output_size

    def _enc(self, dev) -> engine.EncoderPack:
        return self._packed(lambda: engine.EncoderPack(self._sd("encoder."), dev, heads=self.heads, eps=1e-5,
                                                       nparts=self._nparts))

    def forward(self, xs_pad: torch.Tensor, ilens: torch.Tensor):
        _require_cuda(xs_pad, "SenseVoiceEncoderSmall")
        lens = _lengths_list(ilens)
        dev = xs_pad.device
        with torch.cuda.device(dev):
            cu_rows = tables.cu_from_lengths(lens, torch.int32, dev)
            enc = self._enc(dev)
            x = ops.pack_rows(xs_pad.float(), cu_rows, sum(lens))
            x = enc.prologue(x, cu_rows)  # x * sqrt(d) + sinusoidal position code (:561-563)
            y = enc.run(x, cu_rows, max(lens), tp=True)
            out = ops.unpack_rows(y, cu_rows, xs_pad.shape[1])
        olens = torch.as_tensor(lens, dtype=torch.int32, device=dev)  # masks.squeeze(1).sum(1).int()
        return out, olens


class ConformerEncoder(_ParamTree):
    """funasr/models/conformer/encoder.py:285-607 with the options of the aishell recipe
    (examples/aishell/conformer/conf/conformer_12e_6d_2048_256.yaml:16-31): `input_layer="conv2d"`, `rel_pos_type`
    left at its default "legacy" (so rel_pos / rel_selfattn resolve to the legacy classes, encoder.py:352-356),
    swish, macaron FFN pair, convolution module, pre-norm.  `forward(xs_pad, ilens) -> (xs_pad, olens, None)`.

    Every utterance is computed with batch-1 semantics: olens[b] = ((ilens[b] - 1) // 2 - 1) // 2 and its rows do not
    see the padding frames of the batch (the reference's padded run lets the two stride-2 convolutions read them and
    counts ceil-rounded mask lengths for the shorter utterances, subsampling.py:94)."""

    def __init__(self, input_size: int, output_size: int = 256, attention_heads: int = 4, linear_units: int = 2048,
                 num_blocks: int = 6, dropout_rate: float = 0.1, positional_dropout_rate: float = 0.1,
                 attention_dropout_rate: float = 0.0, input_layer: str = "conv2d", normalize_before: bool = True,
                 concat_after: bool = False, positionwise_layer_type: str = "linear",
                 positionwise_conv_kernel_size: int = 3, macaron_style: bool = False, rel_pos_type: str = "legacy",
                 pos_enc_layer_type: str = "rel_pos", selfattention_layer_type: str = "rel_selfattn",
                 activation_type: str = "swish", use_cnn_module: bool = True, zero_triu: bool = False,
                 cnn_module_kernel: int = 31, padding_idx: int = -1, interctc_layer_idx: List[int] = [],
                 interctc_use_conditioning: bool = False, stochastic_depth_rate=0.0, precision: Optional[str] = None,
                 **kwargs):
        super().__init__()
        unsupported = []
        if input_layer != "conv2d":
            unsupported.append("input_layer=%r (only 'conv2d')" % (input_layer,))
        if rel_pos_type != "legacy" or pos_enc_layer_type not in ("rel_pos", "legacy_rel_pos") or \
                selfattention_layer_type not in ("rel_selfattn", "legacy_rel_selfattn"):
            unsupported.append("position / attention types other than the legacy relative ones")
        if not (macaron_style and use_cnn_module and normalize_before) or concat_after or zero_triu:
            unsupported.append("non-macaron / no cnn module / post-norm / concat_after / zero_triu")
        if activation_type != "swish" or positionwise_layer_type != "linear" or interctc_layer_idx:
            unsupported.append("activation other than swish / non-linear positionwise layer / interctc")
        if attention_heads != 4 or output_size // attention_heads != 64:
            unsupported.append("%d heads of dim %d (4 x 64 only)" % (attention_heads, output_size // attention_heads))
        if cnn_module_kernel not in (7, 15, 31):
            unsupported.append("cnn_module_kernel %d (7, 15, 31)" % cnn_module_kernel)
        if unsupported:
            raise NotImplementedError("funasr_b200.ConformerEncoder does not support: " + "; ".join(unsupported))
        self._output_size, self.precision = output_size, precision
        self.interctc_layer_idx, self.interctc_use_conditioning, self.conditioning_layer = [], False, None
        self.normalize_before = True
        self._declare(params.conformer_encoder_shapes(input_size, output_size, linear_units, num_blocks, cnn_module_kernel, ""), "")
        # BatchNorm1d running statistics are buffers in the reference (and in its checkpoints)
        for blk in self._modules["encoders"]._modules.values():
            bn = blk._modules["conv_module"]._modules["norm"]
            for name, fill in (("running_mean", 0.0), ("running_var", 1.0)):
                shape = bn._parameters.pop(name).shape
                bn.register_buffer(name, torch.full(shape, fill))
            bn.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))

    def output_size(self) -> int:
        return self._output_size

    def _state_key(self):
        return tuple((t.data_ptr(), t._version, str(t.device)) for t in list(self.parameters()) + list(self.buffers()))

    def _enc(self, dev) -> "conformer.ConformerEncoderPack":
        return self._packed(lambda: conformer.ConformerEncoderPack(self._sd("encoder."), dev, nparts=self._nparts))

    def forward(self, xs_pad: torch.Tensor, ilens: torch.Tensor, prev_states: torch.Tensor = None, ctc=None):
        _require_cuda(xs_pad, "ConformerEncoder")
        lens = _lengths_list(ilens)
        dev = xs_pad.device
        if xs_pad.shape[1] < 7 or min(lens) < 7:  # check_short_utt / TooShortUttError (encoder.py:555-563)
            raise RuntimeError("has %d frames and is too short for subsampling (it needs more than 7 frames), return empty "
                               "results" % min(min(lens), xs_pad.shape[1]))
        with torch.cuda.device(dev):
            cu_frames = tables.cu_from_lengths(lens, torch.int32, dev)
            enc = self._enc(dev)
            feats = ops.pack_rows(xs_pad.float(), cu_frames, sum(lens))
            x, cu_rows, rows = enc.embed(feats, cu_frames, lens)
            y = enc.run(x, cu_rows, max(rows))
            out = ops.unpack_rows(y, cu_rows, conformer.subsampled_lengths(xs_pad.shape[1])[1])
        olens = torch.as_tensor(rows, dtype=torch.int64, device=dev)
        return out, olens, None


class CifPredictorV2(_ParamTree):
    """funasr/models/paraformer/cif_predictor.py:172-253 at inference (no target labels)."""

    def __init__(self, idim, l_order, r_order, threshold=1.0, dropout=0.1, smooth_factor=1.0, noise_threshold=0,
                 tail_threshold=0.0, tf2torch_tensor_name_prefix_torch="predictor",
                 tf2torch_tensor_name_prefix_tf="seq2seq/cif", tail_mask=True, precision: Optional[str] = None, **kwargs):
        super().__init__()
        if l_order != 1 or r_order != 1:
            raise NotImplementedError("funasr_b200.CifPredictorV2: only l_order = r_order = 1")
        if tail_threshold <= 0.0 or not tail_mask:
            raise NotImplementedError("funasr_b200.CifPredictorV2: tail_threshold > 0 with tail_mask is required")
        self.threshold, self.smooth_factor, self.noise_threshold = threshold, smooth_factor, noise_threshold
        self.tail_threshold, self.tail_mask, self.precision, self.idim = tail_threshold, tail_mask, precision, idim
        self._declare(params.cif_predictor_shapes(idim, l_order, r_order, ""), "")

    def _pred(self, dev) -> engine.PredictorPack:
        return self._packed(lambda: engine.PredictorPack(self._sd("predictor."), dev, nparts=self._nparts,
                                                         smooth_factor=self.smooth_factor,
                                                         noise_threshold=self.noise_threshold,
                                                         tail_threshold=self.tail_threshold))

    def forward(self, hidden, target_label=None, mask=None, ignore_id=-1, mask_chunk_predictor=None,
                target_label_length=None):
        _require_cuda(hidden, "CifPredictorV2")
        if target_label is not None or target_label_length is not None or mask_chunk_predictor is not None:
            raise NotImplementedError("funasr_b200.CifPredictorV2: training-time arguments are not supported")
        b, t, d = hidden.shape
        dev = hidden.device
        lens = [t] * b if mask is None else _lengths_list(mask.reshape(b, -1).sum(-1))
        with torch.cuda.device(dev):
            cu_rows = tables.cu_from_lengths(lens, torch.int32, dev)
            pred = self._pred(dev)
            h = ops.pack_rows(hidden.float(), cu_rows, sum(lens))
            halo = ops.split_rows_halo(h, self._nparts, cu_rows)
            out = pred.run(h, halo, cu_rows, b, want_fires=True)
            tok = out["token_num"]
            lmax = max(tok) if tok else 0
            embeds = ops.unpack_rows(out["embeds"], out["cu_tok"], lmax) if lmax > 0 else hidden.new_zeros(b, 0, d)
            # alphas / cif_peak in the reference's padded [B, T + 1] layout (tail frame at index len_b)
            alphas = hidden.new_zeros(b, t + 1)
            peaks = hidden.new_zeros(b, t + 1)
            a, f = out["alphas"], out["fires"]
            o = 0
            for i, n in enumerate(lens):
                alphas[i, :n] = a[o: o + n]
                alphas[i, n] = self.tail_threshold
                peaks[i, : n + 1] = f[o + i: o + i + n + 1]
                o += n
        token_num = torch.as_tensor(tok, dtype=torch.float32, device=dev)
        return embeds, token_num, alphas, peaks


class CifPredictorV3(CifPredictorV2):
    """funasr/models/bicif_paraformer/cif_predictor.py:96-246 at inference: the V2 head (same tensors; `cif` is the same
    recurrence as `cif_v1`) plus the parameters of the time-stamp head (`upsample_cnn`, `blstm`, `cif_output2`), which
    `BiCifParaformer.inference` runs through `bicif.TimestampPack`.  `forward` returns the reference's 5-tuple; the
    fifth value (`token_num2`, a training loss input) is None."""

    def __init__(self, idim, l_order, r_order, threshold=1.0, dropout=0.1, smooth_factor=1.0, noise_threshold=0,
                 tail_threshold=0.0, tf2torch_tensor_name_prefix_torch="predictor",
                 tf2torch_tensor_name_prefix_tf="seq2seq/cif", smooth_factor2=1.0, noise_threshold2=0, upsample_times=5,
                 upsample_type="cnn", use_cif1_cnn=True, tail_mask=True, precision: Optional[str] = None, **kwargs):
        super().__init__(idim, l_order, r_order, threshold=threshold, dropout=dropout, smooth_factor=smooth_factor,
                         noise_threshold=noise_threshold, tail_threshold=tail_threshold, tail_mask=tail_mask, precision=precision)
        if upsample_type != "cnn_blstm" or use_cif1_cnn or upsample_times != 3:
            raise NotImplementedError("funasr_b200.CifPredictorV3: only upsample_type = cnn_blstm, use_cif1_cnn = false, "
                                      "upsample_times = 3 (funasr/models/bicif_paraformer/template.yaml:47-58)")
        self.smooth_factor2, self.noise_threshold2, self.upsample_times = smooth_factor2, noise_threshold2, upsample_times
        extra = {k: v for k, v in params.bicif_predictor_shapes(idim, upsample_times, "").items() if k not in dict(self.named_parameters())}
        self._declare(extra, "")

    def forward(self, hidden, target_label=None, mask=None, ignore_id=-1, mask_chunk_predictor=None,
                target_label_length=None):
        return super().forward(hidden, target_label, mask, ignore_id, mask_chunk_predictor, target_label_length) + (None,)


class ParaformerSANMDecoder(_ParamTree):
    """funasr/models/paraformer/decoder.py:224-411: `forward(hs_pad, hlens, ys_in_pad, ys_in_lens)`
    -> (logits [B, Lmax, V], olens)."""

    def __init__(self, vocab_size: int, encoder_output_size: int, attention_heads: int = 4, linear_units: int = 2048,
                 num_blocks: int = 6, dropout_rate: float = 0.1, positional_dropout_rate: float = 0.1,
                 self_attention_dropout_rate: float = 0.0, src_attention_dropout_rate: float = 0.0,
                 input_layer: str = "embed", use_output_layer: bool = True, wo_input_layer: bool = False,
                 pos_enc_class=None, normalize_before: bool = True, concat_after: bool = False, att_layer_num: int = 6,
                 kernel_size: int = 21, sanm_shfit: int = 0, lora_list: List[str] = None, lora_rank: int = 8,
                 lora_alpha: int = 16, lora_dropout: float = 0.1, chunk_multiply_factor: tuple = (1,),
                 tf2torch_tensor_name_prefix_torch: str = "decoder",
                 tf2torch_tensor_name_prefix_tf: str = "seq2seq/decoder", precision: Optional[str] = None, **kwargs):
        super().__init__()
        if not normalize_before or concat_after or lora_list or not use_output_layer:
            raise NotImplementedError("funasr_b200.ParaformerSANMDecoder: post-norm / concat_after / lora / no output layer")
        if sanm_shfit not in (0, None):
            raise NotImplementedError("funasr_b200.ParaformerSANMDecoder: sanm_shfit=%r (centred FSMN only)" % (sanm_shfit,))
        if encoder_output_size // attention_heads != 128:
            raise NotImplementedError("funasr_b200.ParaformerSANMDecoder: head dim 128 only")
        self.heads, self.precision, self.vocab_size = attention_heads, precision, vocab_size
        self._declare(params.paraformer_decoder_shapes(vocab_size, encoder_output_size, linear_units, att_layer_num,
                                                       num_blocks, kernel_size, ""), "")

    def _dec(self, dev) -> engine.DecoderPack:
        return self._packed(lambda: engine.DecoderPack(self._sd("decoder."), dev, heads=self.heads, eps=1e-12,
                                                       nparts=self._nparts))

    def forward(self, hs_pad, hlens, ys_in_pad, ys_in_lens, chunk_mask=None, return_hidden: bool = False,
                return_both: bool = False):
        _require_cuda(hs_pad, "ParaformerSANMDecoder")
        if chunk_mask is not None or return_hidden or return_both:
            raise NotImplementedError("funasr_b200.ParaformerSANMDecoder: chunk_mask / return_hidden are not supported")
        dev = hs_pad.device
        rows, toks = _lengths_list(hlens), _lengths_list(ys_in_lens)
        with torch.cuda.device(dev):
            cu_rows = tables.cu_from_lengths(rows, torch.int32, dev)
            cu_tok = tables.cu_from_lengths(toks, torch.int32, dev)
            dec = self._dec(dev)
            enc = ops.pack_rows(hs_pad.float(), cu_rows, sum(rows))
            emb = ops.pack_rows(ys_in_pad.float(), cu_tok, sum(toks))
            _, _, logits = dec.run(enc, cu_rows, emb, cu_tok, max(toks) if toks else 0, want_logits=True)
            out = ops.unpack_rows(logits, cu_tok, ys_in_pad.shape[1])
        return out, torch.as_tensor(toks, dtype=torch.int64, device=dev)


def _as_waveforms(data_in, data_lengths, fs: int, kwargs) -> List["np.ndarray"]:
    """The waveforms of one `inference` call as 1-D float32 host arrays.  Arrays / tensors (what
    `AutoModel.generate(input=ndarray)` and the VAD segment batching of
    funasr/auto/auto_model.py:423-459 pass) are taken as they are; anything else (paths, urls, bytes,
    kaldi ark) goes through the host application's own loader, funasr/utils/load_utils.py:32-111."""
    import numpy as np

    def one(x):
        if isinstance(x, torch.Tensor):
            x = x.detach().cpu().numpy()
        if isinstance(x, np.ndarray):
            return np.ascontiguousarray(x.reshape(-1), dtype=np.float32)
        return None

    if isinstance(data_in, torch.Tensor) and data_in.dim() == 2:
        lens = _lengths_list(data_lengths) if data_lengths is not None else [data_in.shape[1]] * data_in.shape[0]
        return [one(data_in[i, :n]) for i, n in enumerate(lens)]
    items = list(data_in) if isinstance(data_in, (list, tuple)) else [data_in]
    waves = [one(x) for x in items]
    same_rate = int(kwargs.get("fs", 16000)) == int(fs)  # else the loader resamples (load_utils.py:91-96)
    if all(w is not None for w in waves) and same_rate:
        return waves
    try:
        from funasr.utils.load_utils import load_audio_text_image_video  # noqa: WPS433 (host application)
    except ImportError as e:
        raise TypeError("funasr_b200 inference: the input needs the reference's loader (file / url / bytes input, or "
                        "resampling from fs=%s to %s), and funasr.utils.load_utils is not importable"
                        % (kwargs.get("fs", 16000), fs)) from e
    loaded = load_audio_text_image_video(data_in, fs=fs, audio_fs=kwargs.get("fs", 16000),
                                         data_type=kwargs.get("data_type", "sound"), tokenizer=kwargs.get("tokenizer"))
    return [one(x) for x in (loaded if isinstance(loaded, (list, tuple)) else [loaded])]


def _check_frontend_geometry(frontend, who: str) -> None:
    """The packed engines are built for the templates' frontend geometry (16 kHz, 80 mel bins, 25 / 10 ms frames, LFR 7 / 6,
    x 32768 upscaling) and take only the CMVN statistics and the window from the frontend object: anything else must fail
    loudly instead of being silently ignored.  `dither` is NOT honoured by the `inference` overrides -- they behave as
    dither = 0 (deterministic; the reference's default 1.0 adds noise per frame element, funasr/frontends/wav_frontend.py:94)."""
    want = dict(fs=16000, n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, upsacle_samples=True, snip_edges=True)
    for name, value in want.items():
        got = getattr(frontend, name, value)
        if got != value:
            raise NotImplementedError("funasr_b200.%s: frontend.%s = %r is outside the accelerated path (only %r)"
                                      % (who, name, got, value))


class ParaformerInference:
    """`Paraformer.inference` (funasr/models/paraformer/model.py:443-596) as ONE call into the packed
    engine: no padded tensors, no per-utterance device->host reads, no Python loop over frames.  Mixed
    into the reference's own `Paraformer` by `install()` (so construction, checkpoints and training are
    the reference's), or into the stand-alone `Paraformer` below.  Returns the reference's
    `(results, meta_data)`; `[]` when no utterance has a token (model.py:514-515).  Beam search / LM /
    CTC rescoring, fbank inputs and timestamps are not on this path: those calls go to the reference
    method when there is one."""

    _fb200_engine = None
    _fb200_engine_key = None
    _fb200_engine_cls = engine.ParaformerEngine
    _fb200_timestamps = False  # BiCifParaformer: time stamps are part of the path
    _fb200_hotwords = False    # SeacoParaformer: a `hotword` argument leaves the accelerated path
    _fb200_empty = []          # what the reference returns when no utterance has a token

    def _engine_for(self, frontend) -> engine.ParaformerEngine:
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("funasr_b200.Paraformer: the model lives on %s -- this path runs on a CUDA device only "
                               "(no CPU fallback); build the model with device='cuda'" % dev)
        _check_frontend_geometry(frontend, "Paraformer.inference")
        key = (tuple((p.data_ptr(), p._version) for p in self.parameters()), str(dev), PRECISION,
               getattr(frontend, "cmvn_file", None), getattr(frontend, "window", "hamming"))
        if self._fb200_engine is None or self._fb200_engine_key != key:
            sd = {k: v.detach() for k, v in self.state_dict().items()}
            self._fb200_engine = self._fb200_engine_cls(sd, dev, precision=PRECISION, cmvn=getattr(frontend, "cmvn", None),
                                                        window=getattr(frontend, "window", "hamming"))
            if hasattr(self._fb200_engine, "NO_BIAS") and hasattr(self, "NO_BIAS"):
                self._fb200_engine.NO_BIAS = int(self.NO_BIAS)
            self._fb200_engine_key = key
        return self._fb200_engine

    def inference(self, data_in, data_lengths=None, key: list = None, tokenizer=None, frontend=None, **kwargs):
        import time

        off_path = (kwargs.get("decoding_ctc_weight", 0.0) > 0.00001 and getattr(self, "ctc", None) is not None) or (
            kwargs.get("lm_weight", 0.0) > 0.00001 and kwargs.get("lm_file") is not None) or (
            kwargs.get("pred_timestamp", False) and not self._fb200_timestamps) or kwargs.get(
            "data_type", "sound") == "fbank" or getattr(self, "beam_search", None) is not None
        hot_ids = None
        if self._fb200_hotwords and kwargs.get("hotword") is not None and not off_path:
            hot_ids = self._fb200_hotword_ids(kwargs.get("hotword"), tokenizer, frontend)
        if off_path:
            ref = getattr(super(), "inference", None)
            if ref is None:
                raise NotImplementedError("funasr_b200.Paraformer.inference: beam search / LM / CTC rescoring, fbank input "
                                          "and timestamps are outside the accelerated path")
            return ref(data_in, data_lengths=data_lengths, key=key, tokenizer=tokenizer, frontend=frontend, **kwargs)
        meta_data = {}
        t1 = time.perf_counter()
        waves = _as_waveforms(data_in, data_lengths, frontend.fs, dict(kwargs, tokenizer=tokenizer))
        t2 = time.perf_counter()
        meta_data["load_data"] = f"{t2 - t1:0.3f}"
        eng = self._engine_for(frontend)
        stamps = None
        if self._fb200_hotwords:
            ids, stamps, scores = eng.recognize(waves, with_scores=True, begin_time=kwargs.get("begin_time", 0), hotwords=hot_ids)
        elif self._fb200_timestamps:
            ids, stamps, scores = eng.recognize(waves, with_scores=True, begin_time=kwargs.get("begin_time", 0))
        else:
            ids, scores = eng.recognize(waves, with_scores=True)
        t3 = time.perf_counter()
        meta_data["extract_feat"] = f"{t3 - t2:0.3f}"  # here: the whole device path (the reference times fbank alone)
        _, rows = eng.frontend.plan([len(w) for w in waves])
        meta_data["batch_data_time"] = sum(rows) * frontend.frame_shift * frontend.lfr_n / 1000
        if not any(eng.last_token_num):
            return self._fb200_empty
        b = len(waves)
        key = list(key) if key is not None else ["utt%d" % i for i in range(b)]
        if key and isinstance(key[0], (list, tuple)):
            key = list(key[0])
        if len(key) < b:
            key = key * b
        writer = None
        if kwargs.get("output_dir") is not None:
            if not hasattr(self, "writer"):
                from funasr.utils.datadir_writer import DatadirWriter  # noqa: WPS433 (host application)

                self.writer = DatadirWriter(kwargs.get("output_dir"))
            writer = self.writer["1best_recog"]
        results = []
        for i in range(b):
            token_int = ids[i]  # blank / sos / eos already dropped (model.py:561-565)
            if tokenizer is None:
                results.append({"key": key[i], "token_int": token_int})
                continue
            token = tokenizer.ids2tokens(token_int)
            text = tokenizer.tokens2text(token)
            if stamps is not None:  # bicif_paraformer/model.py:343-361: text and stamps post-processed together
                stamp = stamps[i]
                try:
                    from funasr.utils import postprocess_utils  # noqa: WPS433 (host application; text only)

                    text, stamp, _ = postprocess_utils.sentence_postprocess(token, stamp)
                except ImportError:
                    pass
                results.append({"key": key[i], "text": text, "timestamp": stamp})
                if writer is not None:
                    writer["token"][key[i]] = " ".join(token)
                    writer["timestamp"][key[i]] = stamp
                    writer["text"][key[i]] = text
                continue
            if not hasattr(tokenizer, "bpemodel"):
                try:
                    from funasr.utils import postprocess_utils  # noqa: WPS433 (host application; text only)

                    text, _ = postprocess_utils.sentence_postprocess(token)
                except ImportError:
                    pass
            results.append({"key": key[i], "text": text})
            if writer is not None:
                writer["token"][key[i]] = " ".join(token)
                writer["text"][key[i]] = text
        return results, meta_data


class _CTCHead(_ParamTree):
    """`ctc.ctc_lo` of funasr/models/ctc/ctc.py:27-33: present in checkpoints trained with ctc_weight > 0."""

    def __init__(self, odim: int, encoder_output_size: int, **kwargs):
        super().__init__()
        self._declare({"ctc_lo.weight": (odim, encoder_output_size), "ctc_lo.bias": (odim,)}, "")


class Paraformer(ParaformerInference, nn.Module):
    """Stand-alone `Paraformer` (funasr/models/paraformer/model.py:30-175) for hosts without the reference
    package: same constructor keywords, sub-module names (`encoder`, `predictor`, `decoder`, `ctc`) and
    therefore state_dict keys; inference only (`forward`, the training loss, is the reference's business)."""

    def __init__(self, specaug=None, specaug_conf=None, normalize=None, normalize_conf=None, encoder: str = None,
                 encoder_conf: Optional[Dict] = None, decoder: str = None, decoder_conf: Optional[Dict] = None,
                 ctc: str = None, ctc_conf: Optional[Dict] = None, predictor: str = None,
                 predictor_conf: Optional[Dict] = None, ctc_weight: float = 0.5, input_size: int = 80,
                 vocab_size: int = -1, ignore_id: int = -1, blank_id: int = 0, sos: int = 1, eos: int = 2,
                 lsm_weight: float = 0.0, length_normalized_loss: bool = False, predictor_weight: float = 0.0,
                 predictor_bias: int = 0, sampling_ratio: float = 0.2, share_embedding: bool = False,
                 use_1st_decoder_loss: bool = False, **kwargs):
        super().__init__()
        if share_embedding:
            raise NotImplementedError("funasr_b200.Paraformer: share_embedding is a training-time option")

        def lookup(table, name):
            cls = DROPIN_CLASSES.get((table, name))
            if cls is None:
                raise NotImplementedError("funasr_b200.Paraformer: no drop-in for %s[%s]" % (table, name))
            return cls

        self.encoder = lookup("encoder_classes", encoder)(input_size=input_size, **(encoder_conf or {}))
        self.decoder = None
        if decoder is not None and ctc_weight != 1.0:
            self.decoder = lookup("decoder_classes", decoder)(vocab_size=vocab_size, encoder_output_size=self.encoder.output_size(),
                                                              **(decoder_conf or {}))
        self.ctc = _CTCHead(vocab_size, self.encoder.output_size(), **(ctc_conf or {})) if ctc_weight > 0.0 else None
        self.predictor = lookup("predictor_classes", predictor)(**(predictor_conf or {}))
        self.blank_id, self.vocab_size, self.ignore_id, self.ctc_weight = blank_id, vocab_size, ignore_id, ctc_weight
        self.sos = sos if sos is not None else vocab_size - 1
        self.eos = eos if eos is not None else vocab_size - 1
        self.predictor_weight, self.predictor_bias, self.sampling_ratio = predictor_weight, predictor_bias, sampling_ratio
        self.beam_search = None

    def forward(self, *args, **kwargs):
        raise NotImplementedError("funasr_b200.Paraformer: the training forward is not part of the accelerated path")


class SenseVoiceInference:
    """`SenseVoiceSmall.inference` (funasr/models/sense_voice/model.py:809-906) as one call into the packed
    engine: query rows, 50 + 20 SAN-M layers, CTC projection, per-frame argmax, unique_consecutive and blank
    removal all stay on the device; the host gets token ids.  Returns the reference's `(results, meta_data)`.
    `output_timestamp` (forced alignment) and fbank inputs go to the reference method when there is one."""

    _fb200_engine = None
    _fb200_engine_key = None
    LID = engine.SenseVoiceEngine.LID
    TEXTNORM = engine.SenseVoiceEngine.TEXTNORM
    EMO_UNK = 25009  # emo_dict["unk"], model.py:649-655

    def _engine_for(self, frontend) -> engine.SenseVoiceEngine:
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("funasr_b200.SenseVoiceSmall: the model lives on %s -- this path runs on a CUDA device "
                               "only (no CPU fallback); build the model with device='cuda'" % dev)
        _check_frontend_geometry(frontend, "SenseVoiceSmall.inference")
        key = (tuple((p.data_ptr(), p._version) for p in self.parameters()), str(dev), PRECISION,
               getattr(frontend, "cmvn_file", None), getattr(frontend, "window", "hamming"))
        if self._fb200_engine is None or self._fb200_engine_key != key:
            sd = {k: v.detach() for k, v in self.state_dict().items()}
            self._fb200_engine = engine.SenseVoiceEngine(sd, dev, precision=PRECISION, cmvn=getattr(frontend, "cmvn", None),
                                                         window=getattr(frontend, "window", "hamming"),
                                                         blank_id=getattr(self, "blank_id", 0))
            self._fb200_engine_key = key
        return self._fb200_engine

    def inference(self, data_in, data_lengths=None, key: list = ("wav_file_tmp_name",), tokenizer=None, frontend=None,
                  **kwargs):
        import time

        if kwargs.get("output_timestamp", False) or kwargs.get("data_type", "sound") == "fbank":
            ref = getattr(super(), "inference", None)
            if ref is None:
                raise NotImplementedError("funasr_b200.SenseVoiceSmall.inference: timestamps (forced alignment) and "
                                          "fbank input are outside the accelerated path")
            return ref(data_in, data_lengths=data_lengths, key=key, tokenizer=tokenizer, frontend=frontend, **kwargs)
        meta_data = {}
        t1 = time.perf_counter()
        waves = _as_waveforms(data_in, data_lengths, frontend.fs, dict(kwargs, tokenizer=tokenizer))
        t2 = time.perf_counter()
        meta_data["load_data"] = f"{t2 - t1:0.3f}"
        language = kwargs.get("language", "auto")
        language = language if language in self.LID else "auto"  # unknown names fall back to id 0 (model.py:853-855)
        textnorm = kwargs.get("text_norm", None)
        if textnorm is None:
            textnorm = "withitn" if kwargs.get("use_itn", False) else "woitn"
        if textnorm not in self.TEXTNORM:
            raise KeyError(textnorm)
        eng = self._engine_for(frontend)
        ban = (self.EMO_UNK,) if kwargs.get("ban_emo_unk", False) else ()
        ids = eng.recognize(waves, language=language, textnorm=textnorm, ban_ids=ban)
        t3 = time.perf_counter()
        meta_data["extract_feat"] = f"{t3 - t2:0.3f}"  # here: the whole device path
        _, rows = eng.frontend.plan([len(w) for w in waves])
        meta_data["batch_data_time"] = sum(rows) * frontend.frame_shift * frontend.lfr_n / 1000
        b = len(waves)
        key = list(key)
        if key and isinstance(key[0], (list, tuple)):
            key = list(key[0])
        if len(key) < b:
            key = key * b
        writer = None
        if kwargs.get("output_dir") is not None:
            if not hasattr(self, "writer"):
                from funasr.utils.datadir_writer import DatadirWriter  # noqa: WPS433 (host application)

                self.writer = DatadirWriter(kwargs.get("output_dir"))
            writer = self.writer["1best_recog"]
        results = []
        for i in range(b):
            if tokenizer is None:
                results.append({"key": key[i], "token_int": ids[i]})
                continue
            text = tokenizer.decode(ids[i])
            if writer is not None:
                writer["text"][key[i]] = text
            results.append({"key": key[i], "text": text})
        return results, meta_data


class SenseVoiceSmall(SenseVoiceInference, nn.Module):
    """Stand-alone `SenseVoiceSmall` (funasr/models/sense_voice/model.py:588-660) for hosts without the
    reference package: same constructor keywords and state_dict keys (`encoder.*`, `ctc.ctc_lo.*`,
    `embed.weight`); inference only."""

    def __init__(self, specaug=None, specaug_conf=None, normalize=None, normalize_conf=None, encoder: str = None,
                 encoder_conf: Optional[Dict] = None, ctc_conf: Optional[Dict] = None, input_size: int = 80,
                 vocab_size: int = -1, ignore_id: int = -1, blank_id: int = 0, sos: int = 1, eos: int = 2,
                 length_normalized_loss: bool = False, **kwargs):
        super().__init__()
        cls = DROPIN_CLASSES.get(("encoder_classes", encoder))
        if cls is None:
            raise NotImplementedError("funasr_b200.SenseVoiceSmall: no drop-in for encoder_classes[%s]" % encoder)
        self.encoder = cls(input_size=input_size, **(encoder_conf or {}))
        self.encoder_output_size = self.encoder.output_size()
        self.ctc = _CTCHead(vocab_size, self.encoder_output_size, **(ctc_conf or {}))
        self.blank_id, self.vocab_size, self.ignore_id = blank_id, vocab_size, ignore_id
        self.sos = sos if sos is not None else vocab_size - 1
        self.eos = eos if eos is not None else vocab_size - 1
        self.lid_dict, self.textnorm_dict = dict(self.LID), dict(self.TEXTNORM)
        self.emo_dict = {"unk": 25009, "happy": 25001, "sad": 25002, "angry": 25003, "neutral": 25004}
        self.embed = nn.Embedding(7 + len(self.lid_dict) + len(self.textnorm_dict), input_size)

    def forward(self, *args, **kwargs):
        raise NotImplementedError("funasr_b200.SenseVoiceSmall: the training forward is not part of the accelerated path")


DROPIN_CLASSES = {
    ("frontend_classes", "WavFrontend"): WavFrontend,
    ("frontend_classes", "wav_frontend"): WavFrontend,
    ("encoder_classes", "SANMEncoder"): SANMEncoder,
    ("encoder_classes", "SenseVoiceEncoderSmall"): SenseVoiceEncoderSmall,
    ("predictor_classes", "CifPredictorV2"): CifPredictorV2,
    ("decoder_classes", "ParaformerSANMDecoder"): ParaformerSANMDecoder,
    ("model_classes", "Paraformer"): Paraformer,
    ("model_classes", "SenseVoiceSmall"): SenseVoiceSmall,
    ("encoder_classes", "ConformerEncoder"): ConformerEncoder,
}


class CTCInference:
    """`inference` of `model_classes["CTC"]` (funasr/models/ctc/model.py:162-247) over a ConformerEncoder as ONE
    call into the packed engine (BASELINE config 4): waveforms -> fbank -> conv2d subsampling -> Conformer blocks ->
    ctc_lo -> per-frame argmax -> unique_consecutive -> blank / sos / eos removed.  Returns the reference's
    `(results, meta_data)` with `{"key", "token", "text"}` entries.  Unlike the reference method (which raises for
    batch_size > 1, :172-173) a list of inputs is decoded as one packed batch.  Other encoders and fbank inputs go
    to the reference method when there is one."""

    _fb200_engine = None
    _fb200_engine_key = None

    def _engine_for(self, frontend) -> "conformer.ConformerCTCEngine":
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("funasr_b200.CTC: the model lives on %s -- this path runs on a CUDA device only "
                               "(no CPU fallback); build the model with device='cuda'" % dev)
        key = (tuple((t.data_ptr(), t._version) for t in list(self.parameters()) + list(self.buffers())), str(dev), PRECISION,
               getattr(frontend, "window", "hamming"))
        if self._fb200_engine is None or self._fb200_engine_key != key:
            sd = {k: v.detach() for k, v in self.state_dict().items()}
            prec = "fp32" if PRECISION == "fp32x6" else PRECISION
            self._fb200_engine = conformer.ConformerCTCEngine(sd, dev, precision=prec, window=getattr(frontend, "window", "hamming"))
            self._fb200_engine_key = key
        return self._fb200_engine

    def inference(self, data_in, data_lengths=None, key: list = None, tokenizer=None, frontend=None, **kwargs):
        import time

        enc = getattr(self, "encoder", None)
        on_path = "encoder.embed.conv.0.weight" in self.state_dict() and "encoder.encoders.0.conv_module.norm.running_var" in \
            self.state_dict() and kwargs.get("data_type", "sound") != "fbank" and getattr(frontend, "lfr_m", 1) == 1 and \
            getattr(frontend, "lfr_n", 1) == 1 and getattr(frontend, "cmvn_file", None) is None
        if not on_path:
            ref = getattr(super(), "inference", None)
            if ref is None:
                raise NotImplementedError("funasr_b200.CTC.inference: only a ConformerEncoder (conv2d input layer) fed by "
                                          "WavFrontend(lfr_m=1, lfr_n=1, no CMVN) is on the accelerated path (encoder: %s)"
                                          % type(enc).__name__)
            return ref(data_in, data_lengths=data_lengths, key=key, tokenizer=tokenizer, frontend=frontend, **kwargs)
        meta_data = {}
        t1 = time.perf_counter()
        waves = _as_waveforms(data_in, data_lengths, frontend.fs, dict(kwargs, tokenizer=tokenizer))
        t2 = time.perf_counter()
        meta_data["load_data"] = f"{t2 - t1:0.3f}"
        eng = self._engine_for(frontend)
        ids = eng.recognize(waves)
        t3 = time.perf_counter()
        meta_data["extract_feat"] = f"{t3 - t2:0.3f}"  # here: the whole device path
        frames = [tables.num_frames(len(w), eng.frontend.frame_len, eng.frontend.frame_shift) for w in waves]
        meta_data["batch_data_time"] = sum(frames) * frontend.frame_shift * frontend.lfr_n / 1000
        b = len(waves)
        key = list(key) if key is not None else ["utt%d" % i for i in range(b)]
        if key and isinstance(key[0], (list, tuple)):
            key = list(key[0])
        if len(key) < b:
            key = key * b
        results = []
        for i in range(b):
            token_int = ids[i]
            if tokenizer is None:
                results.append({"key": key[i], "token_int": token_int})
                continue
            token = tokenizer.ids2tokens(token_int)
            text = tokenizer.tokens2text(token)
            try:
                from funasr.utils import postprocess_utils  # noqa: WPS433 (host application; text only)

                text, _ = postprocess_utils.sentence_postprocess(token)
            except ImportError:
                pass
            results.append({"key": key[i], "token": token, "text": text})
        return results, meta_data


class CTC(CTCInference, nn.Module):
    """Stand-alone `model_classes["CTC"]` (funasr/models/ctc/model.py:17-72) for hosts without the reference package:
    encoder by registry key + `ctc.ctc_lo`, the reference's state_dict keys."""

    def __init__(self, specaug=None, specaug_conf=None, normalize=None, normalize_conf=None, encoder: str = None,
                 encoder_conf: dict = None, ctc_conf: dict = None, input_size: int = 80, vocab_size: int = -1,
                 ignore_id: int = -1, blank_id: int = 0, sos: int = 1, eos: int = 2, length_normalized_loss: bool = False,
                 **kwargs):
        super().__init__()
        cls = DROPIN_CLASSES.get(("encoder_classes", encoder))
        if cls is None:
            raise NotImplementedError("funasr_b200.CTC: encoder %r has no B200 implementation" % (encoder,))
        self.encoder = cls(input_size=input_size, **(encoder_conf or {}))
        self.ctc = _CTCHead(vocab_size, self.encoder.output_size())
        self.blank_id, self.sos, self.eos, self.vocab_size, self.ignore_id = blank_id, sos, eos, vocab_size, ignore_id

    def forward(self, *args, **kwargs):
        raise NotImplementedError("funasr_b200.CTC: training / loss computation is outside the accelerated path")


DROPIN_CLASSES[("model_classes", "CTC")] = CTC


class BiCifParaformerInference(ParaformerInference):
    """`BiCifParaformer.inference` (funasr/models/bicif_paraformer/model.py:222-369) as one call into
    `bicif.BiCifParaformerEngine`: token ids as for Paraformer plus `"timestamp"`: [[begin ms, end ms], ...] per token
    from the CifPredictorV3 time-stamp head (`begin_time` kwarg = the VAD offset, :349).  Returns the reference's
    `(results, meta_data)`."""

    _fb200_timestamps = True

    @property
    def _fb200_engine_cls(self):
        from . import bicif

        return bicif.BiCifParaformerEngine


class BiCifParaformer(BiCifParaformerInference, Paraformer):
    """Stand-alone `BiCifParaformer` (funasr/models/bicif_paraformer/model.py:37-52): Paraformer's constructor with a
    CifPredictorV3 predictor; inference only."""


class SeacoParaformerInference(BiCifParaformerInference):
    """`SeacoParaformer.inference` (funasr/models/seaco_paraformer/model.py:361-510; the class the hub alias
    `paraformer-zh` resolves to) WITHOUT hot words: `_seaco_decode_with_ASF` then returns the plain decoder's log-softmax
    (:252-254, :325-327) and the rest is BiCifParaformer's inference -- pinned by tests/golden/seaco.npz (the unmodified
    reference's SeACo results equal its BiCif results on the same weights).  WITH `hotword=` the hot-word branch runs in
    `seaco.SeacoParaformerEngine` (bias-encoder LSTM, attention-score filter above 50 hot words, SeACo decoder, merge)."""

    _fb200_hotwords = True
    _fb200_empty = ([],)  # model.py:417-418

    @property
    def _fb200_engine_cls(self):
        from . import seaco

        return seaco.SeacoParaformerEngine

    def _fb200_hotword_ids(self, hotword, tokenizer, frontend):
        """Hot words as token-id lists, the last one [sos]: the reference's `generate_hotwords_list` (model.py:512-617:
        string, .txt file or URL, optional seg_dict) when this class wraps the reference model; stand-alone: a list of
        id lists, or a blank-separated string looked up in the tokenizer."""
        gen = getattr(super(), "generate_hotwords_list", None)
        if gen is not None:
            return gen(hotword, tokenizer=tokenizer, frontend=frontend)
        if isinstance(hotword, str):
            if tokenizer is None:
                raise RuntimeError("funasr_b200.SeacoParaformer: a hot-word string needs the tokenizer")
            return [tokenizer.tokens2ids([w]) for w in hotword.strip().split()] + [[self.sos]]
        return [list(map(int, w)) for w in hotword] + [[self.sos]]


class _HotwordTensors(_ParamTree):
    """`bias_encoder.*`, `seaco_decoder.*`, `hotword_output_layer.*` under the reference's names (model.py:66-96)."""

    def __init__(self, shapes):
        super().__init__()
        self._declare(shapes, "")


class SeacoParaformer(SeacoParaformerInference, Paraformer):
    """Stand-alone `SeacoParaformer` (funasr/models/seaco_paraformer/model.py:44-105) for hosts without the reference
    package: BiCifParaformer's modules plus the hot-word tensors under the reference's state_dict keys; inference only."""

    def __init__(self, *args, inner_dim=512, bias_encoder_type="lstm", bias_encoder_bid=False, bias_encoder_dropout_rate=0.0,
                 seaco_lsm_weight=0.0, seaco_length_normal=True, seaco_length_normalized_loss=True, train_decoder=True,
                 NO_BIAS=8377, seaco_decoder=None, seaco_decoder_conf=None, seaco_weight=0.01, **kwargs):
        super().__init__(*args, **kwargs)
        if bias_encoder_type != "lstm" or bias_encoder_bid:
            raise NotImplementedError("funasr_b200.SeacoParaformer: only a unidirectional LSTM bias encoder")
        conf = dict(seaco_decoder_conf or {})
        blocks = max(int(conf.get("att_layer_num", 6)), int(conf.get("num_blocks", 6)))
        extra = {k: v for k, v in params.seaco_paraformer_shapes(self.vocab_size, 1, 1, inner_dim, blocks,
                                                                 int(conf.get("linear_units", 2048)),
                                                                 int(conf.get("kernel_size", 21))).items()
                 if k.startswith(("bias_encoder.", "seaco_decoder.", "hotword_output_layer."))}
        groups = {}
        for k, v in extra.items():
            head, rest = k.split(".", 1)
            groups.setdefault(head, {})[rest] = v
        for head, shapes in groups.items():
            self.add_module(head, _HotwordTensors(shapes))
        self.NO_BIAS = NO_BIAS


DROPIN_CLASSES[("model_classes", "SeacoParaformer")] = SeacoParaformer
DROPIN_CLASSES[("predictor_classes", "CifPredictorV3")] = CifPredictorV3
DROPIN_CLASSES[("model_classes", "BiCifParaformer")] = BiCifParaformer


def _model_class(key: str, cls):
    """The class registered under model_classes[key]: the reference's own class with `inference` replaced
    when the reference is importable (construction, checkpoints and training stay the reference's), else the
    stand-alone class."""
    if cls is Paraformer:
        try:
            from funasr.models.paraformer.model import Paraformer as ref_cls  # noqa: WPS433
        except ImportError:
            return cls
        return type("Paraformer", (ParaformerInference, ref_cls), {"__doc__": ParaformerInference.__doc__,
                                                                    "__module__": __name__})
    if cls is SenseVoiceSmall:
        try:
            from funasr.models.sense_voice.model import SenseVoiceSmall as ref_sv  # noqa: WPS433
        except ImportError:
            return cls
        return type("SenseVoiceSmall", (SenseVoiceInference, ref_sv), {"__doc__": SenseVoiceInference.__doc__,
                                                                        "__module__": __name__})
    if cls is SeacoParaformer:
        try:
            from funasr.models.seaco_paraformer.model import SeacoParaformer as ref_seaco  # noqa: WPS433
        except ImportError:
            return cls
        return type("SeacoParaformer", (SeacoParaformerInference, ref_seaco), {"__doc__": SeacoParaformerInference.__doc__,
                                                                                "__module__": __name__})
    if cls is BiCifParaformer:
        try:
            from funasr.models.bicif_paraformer.model import BiCifParaformer as ref_bicif  # noqa: WPS433
        except ImportError:
            return cls
        return type("BiCifParaformer", (BiCifParaformerInference, ref_bicif), {"__doc__": BiCifParaformerInference.__doc__,
                                                                                "__module__": __name__})
    if cls is CTC:
        try:
            from funasr.models.ctc.model import Transformer as ref_ctc  # noqa: WPS433 (registered as "CTC", model.py:17)
        except ImportError:
            return cls
        return type("CTC", (CTCInference, ref_ctc), {"__doc__": CTCInference.__doc__, "__module__": __name__})
    return cls


def install(tables_obj=None) -> List[str]:
    """Re-register the drop-in classes under the reference's keys (`import funasr` must have run, so
    the reference's own registrations are already in place and ours come last)."""
    if tables_obj is None:
        from funasr.register import tables as tables_obj  # noqa: WPS433 (the host application's registry)
    done = []
    for (table, key), cls in DROPIN_CLASSES.items():
        getattr(tables_obj, table)[key] = _model_class(key, cls) if table == "model_classes" else cls
        done.append("%s[%s]" % (table, key))
    return done
