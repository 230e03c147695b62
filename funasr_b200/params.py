"""Parameter inventory of the models on the path, under the reference's state_dict keys.

The drop-in modules must load a reference checkpoint with `strict=True`, so names and shapes
are the contract ("Parameter names/shapes the drop-in must keep").
The same table drives seeded random init (`synth.synth_state_dict`) on both sides of a
parity check.
"""
from collections import OrderedDict
from typing import Dict, Tuple

Shapes = Dict[str, Tuple[int, ...]]


def _norm(s: Shapes, p: str, d: int) -> None:
    s[p + ".weight"] = (d,)
    s[p + ".bias"] = (d,)


def _linear(s: Shapes, p: str, din: int, dout: int, bias: bool = True) -> None:
    s[p + ".weight"] = (dout, din)
    if bias:
        s[p + ".bias"] = (dout,)


def sanm_encoder_layer_shapes(s: Shapes, p: str, din: int, d: int, ff: int, k: int) -> None:
    """EncoderLayerSANM (funasr/models/sanm/encoder.py:44-70) + MultiHeadedAttentionSANM
    (funasr/models/sanm/attention.py:150-205) + PositionwiseFeedForward."""
    _linear(s, p + "self_attn.linear_out", d, d)
    _linear(s, p + "self_attn.linear_q_k_v", din, 3 * d)
    s[p + "self_attn.fsmn_block.weight"] = (d, 1, k)
    _linear(s, p + "feed_forward.w_1", d, ff)
    _linear(s, p + "feed_forward.w_2", ff, d)
    _norm(s, p + "norm1", din)
    _norm(s, p + "norm2", d)


def sanm_encoder_shapes(input_size: int = 560, output_size: int = 512, linear_units: int = 2048,
                        num_blocks: int = 50, kernel_size: int = 11, prefix: str = "encoder.",
                        tp_blocks: int = 0) -> Shapes:
    """SANMEncoder (funasr/models/sanm/encoder.py:170-360); `tp_blocks` > 0 adds the
    SenseVoiceEncoderSmall tail (funasr/models/sense_voice/model.py:444-551)."""
    s: Shapes = OrderedDict()
    sanm_encoder_layer_shapes(s, prefix + "encoders0.0.", input_size, output_size, linear_units, kernel_size)
    for i in range(num_blocks - 1):
        sanm_encoder_layer_shapes(s, prefix + "encoders.%d." % i, output_size, output_size, linear_units, kernel_size)
    _norm(s, prefix + "after_norm", output_size)
    for i in range(tp_blocks):
        sanm_encoder_layer_shapes(s, prefix + "tp_encoders.%d." % i, output_size, output_size, linear_units, kernel_size)
    if tp_blocks:
        _norm(s, prefix + "tp_norm", output_size)
    return s


def cif_predictor_shapes(idim: int = 512, l_order: int = 1, r_order: int = 1, prefix: str = "predictor.") -> Shapes:
    """CifPredictorV2 (funasr/models/paraformer/cif_predictor.py:172-200)."""
    s: Shapes = OrderedDict()
    s[prefix + "cif_conv1d.weight"] = (idim, idim, l_order + r_order + 1)
    s[prefix + "cif_conv1d.bias"] = (idim,)
    _linear(s, prefix + "cif_output", idim, 1)
    return s


def paraformer_decoder_shapes(vocab_size: int = 8404, d: int = 512, linear_units: int = 2048,
                              att_layer_num: int = 16, num_blocks: int = 16, kernel_size: int = 11,
                              prefix: str = "decoder.") -> Shapes:
    """ParaformerSANMDecoder (funasr/models/paraformer/decoder.py:224-357)."""
    s: Shapes = OrderedDict()
    s[prefix + "embed.0.weight"] = (vocab_size, d)
    _norm(s, prefix + "after_norm", d)
    _linear(s, prefix + "output_layer", d, vocab_size)

    def ffn(p):
        _linear(s, p + "feed_forward.w_1", d, linear_units)
        _linear(s, p + "feed_forward.w_2", linear_units, d, bias=False)
        _norm(s, p + "feed_forward.norm", linear_units)

    for i in range(att_layer_num):
        p = prefix + "decoders.%d." % i
        s[p + "self_attn.fsmn_block.weight"] = (d, 1, kernel_size)
        _linear(s, p + "src_attn.linear_q", d, d)
        _linear(s, p + "src_attn.linear_k_v", d, 2 * d)
        _linear(s, p + "src_attn.linear_out", d, d)
        ffn(p)
        _norm(s, p + "norm1", d)
        _norm(s, p + "norm2", d)
        _norm(s, p + "norm3", d)
    for i in range(max(num_blocks - att_layer_num, 0)):
        p = prefix + "decoders2.%d." % i
        s[p + "self_attn.fsmn_block.weight"] = (d, 1, kernel_size)
        ffn(p)
        _norm(s, p + "norm1", d)
        _norm(s, p + "norm2", d)
    p = prefix + "decoders3.0."
    ffn(p)
    _norm(s, p + "norm1", d)
    return s


def paraformer_large_shapes(vocab_size: int = 8404, enc_blocks: int = 50, dec_blocks: int = 16) -> Shapes:
    """Paraformer-large (funasr/models/paraformer/template.yaml): 220.08 M parameters."""
    s: Shapes = OrderedDict()
    s.update(sanm_encoder_shapes(num_blocks=enc_blocks))
    s.update(paraformer_decoder_shapes(vocab_size, att_layer_num=dec_blocks, num_blocks=dec_blocks))
    s.update(cif_predictor_shapes())
    return s


def sense_voice_small_shapes(vocab_size: int = 25055, enc_blocks: int = 50, tp_blocks: int = 20) -> Shapes:
    """SenseVoiceSmall (funasr/models/sense_voice/model.py:598-660): SenseVoiceEncoderSmall, the CTC
    projection (funasr/models/ctc/ctc.py:28) and the 16 x 560 query embedding (:646-648)."""
    s = sanm_encoder_shapes(num_blocks=enc_blocks, tp_blocks=tp_blocks)
    s["ctc.ctc_lo.weight"] = (vocab_size, 512)
    s["ctc.ctc_lo.bias"] = (vocab_size,)
    s["embed.weight"] = (16, 560)
    return s


def count_params(shapes: Shapes) -> int:
    n = 0
    for shp in shapes.values():
        c = 1
        for d in shp:
            c *= d
        n += c
    return n
