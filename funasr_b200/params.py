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


def bicif_predictor_shapes(idim: int = 512, upsample_times: int = 3, prefix: str = "predictor.") -> Shapes:
    """CifPredictorV3 with `upsample_type: cnn_blstm` (funasr/models/bicif_paraformer/cif_predictor.py:96-169;
    funasr/models/bicif_paraformer/template.yaml:47-58): the V2 tensors plus the upsampling ConvTranspose1d, the
    one-layer bidirectional LSTM and the second alpha head."""
    s = cif_predictor_shapes(idim, prefix=prefix)
    s[prefix + "upsample_cnn.weight"] = (idim, idim, upsample_times)
    s[prefix + "upsample_cnn.bias"] = (idim,)
    for suffix in ("", "_reverse"):
        s[prefix + "blstm.weight_ih_l0" + suffix] = (4 * idim, idim)
        s[prefix + "blstm.weight_hh_l0" + suffix] = (4 * idim, idim)
        s[prefix + "blstm.bias_ih_l0" + suffix] = (4 * idim,)
        s[prefix + "blstm.bias_hh_l0" + suffix] = (4 * idim,)
    _linear(s, prefix + "cif_output2", 2 * idim, 1)
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


def bicif_paraformer_shapes(vocab_size: int = 8404, enc_blocks: int = 50, dec_blocks: int = 16) -> Shapes:
    """BiCifParaformer (funasr/models/bicif_paraformer/template.yaml): Paraformer-large with CifPredictorV3."""
    s: Shapes = OrderedDict()
    s.update(sanm_encoder_shapes(num_blocks=enc_blocks))
    s.update(paraformer_decoder_shapes(vocab_size, att_layer_num=dec_blocks, num_blocks=dec_blocks))
    s.update(bicif_predictor_shapes())
    return s


def seaco_paraformer_shapes(vocab_size: int = 8404, enc_blocks: int = 50, dec_blocks: int = 16, inner_dim: int = 512,
                            seaco_layers: int = 6, seaco_units: int = 1024, seaco_kernel: int = 21) -> Shapes:
    """SeacoParaformer (funasr/models/seaco_paraformer/template.yaml): BiCifParaformer plus the 2-layer bias-encoder LSTM
    (model.py:66-74), the SeACo decoder (a ParaformerSANMDecoder with `wo_input_layer`, no output layer; `num_blocks: 4`
    leaves `att_layer_num` at its default 6, so it has 6 attention layers + the FFN-only layer, decoder.py:303-351) and
    `hotword_output_layer` (:96)."""
    s = bicif_paraformer_shapes(vocab_size, enc_blocks, dec_blocks)
    for layer in range(2):
        s["bias_encoder.weight_ih_l%d" % layer] = (4 * inner_dim, inner_dim)
        s["bias_encoder.weight_hh_l%d" % layer] = (4 * inner_dim, inner_dim)
        s["bias_encoder.bias_ih_l%d" % layer] = (4 * inner_dim,)
        s["bias_encoder.bias_hh_l%d" % layer] = (4 * inner_dim,)
    dec = paraformer_decoder_shapes(vocab_size, inner_dim, seaco_units, seaco_layers, seaco_layers, seaco_kernel, prefix="seaco_decoder.")
    for k, v in dec.items():
        if not k.startswith(("seaco_decoder.embed.", "seaco_decoder.output_layer.")):
            s[k] = v
    _linear(s, "hotword_output_layer", inner_dim, vocab_size)
    return s


def sense_voice_small_shapes(vocab_size: int = 25055, enc_blocks: int = 50, tp_blocks: int = 20) -> Shapes:
    """SenseVoiceSmall (funasr/models/sense_voice/model.py:598-660): SenseVoiceEncoderSmall, the CTC
    projection (funasr/models/ctc/ctc.py:28) and the 16 x 560 query embedding (:646-648)."""
    s = sanm_encoder_shapes(num_blocks=enc_blocks, tp_blocks=tp_blocks)
    s["ctc.ctc_lo.weight"] = (vocab_size, 512)
    s["ctc.ctc_lo.bias"] = (vocab_size,)
    s["embed.weight"] = (16, 560)
    return s


def conformer_encoder_shapes(input_size: int = 80, d: int = 256, linear_units: int = 2048, num_blocks: int = 12,
                             cnn_kernel: int = 15, prefix: str = "encoder.") -> Shapes:
    """ConformerEncoder (funasr/models/conformer/encoder.py:285-528) with the aishell recipe's options
    (examples/aishell/conformer/conf/conformer_12e_6d_2048_256.yaml:16-31): conv2d input layer
    (funasr/models/transformer/utils/subsampling.py:49-94), legacy rel-pos attention
    (funasr/models/transformer/attention.py:221-312), macaron FFN pair, ConvolutionModule (encoder.py:53-120)."""
    s: Shapes = OrderedDict()
    f2 = ((input_size - 1) // 2 - 1) // 2
    s[prefix + "embed.conv.0.weight"] = (d, 1, 3, 3)
    s[prefix + "embed.conv.0.bias"] = (d,)
    s[prefix + "embed.conv.2.weight"] = (d, d, 3, 3)
    s[prefix + "embed.conv.2.bias"] = (d,)
    _linear(s, prefix + "embed.out.0", d * f2, d)
    for i in range(num_blocks):
        p = prefix + "encoders.%d." % i
        for name in ("linear_q", "linear_k", "linear_v", "linear_out"):
            _linear(s, p + "self_attn." + name, d, d)
        _linear(s, p + "self_attn.linear_pos", d, d, bias=False)
        s[p + "self_attn.pos_bias_u"] = (4, d // 4)
        s[p + "self_attn.pos_bias_v"] = (4, d // 4)
        for ff in ("feed_forward", "feed_forward_macaron"):
            _linear(s, p + ff + ".w_1", d, linear_units)
            _linear(s, p + ff + ".w_2", linear_units, d)
        s[p + "conv_module.pointwise_conv1.weight"] = (2 * d, d, 1)
        s[p + "conv_module.pointwise_conv1.bias"] = (2 * d,)
        s[p + "conv_module.depthwise_conv.weight"] = (d, 1, cnn_kernel)
        s[p + "conv_module.depthwise_conv.bias"] = (d,)
        _norm(s, p + "conv_module.norm", d)
        s[p + "conv_module.norm.running_mean"] = (d,)
        s[p + "conv_module.norm.running_var"] = (d,)
        s[p + "conv_module.pointwise_conv2.weight"] = (d, d, 1)
        s[p + "conv_module.pointwise_conv2.bias"] = (d,)
        for name in ("norm_ff", "norm_mha", "norm_ff_macaron", "norm_conv", "norm_final"):
            _norm(s, p + name, d)
    _norm(s, prefix + "after_norm", d)
    return s


def conformer_ctc_shapes(vocab_size: int = 4234, num_blocks: int = 12, d: int = 256) -> Shapes:
    """`model_classes["CTC"]` (funasr/models/ctc/model.py:17-72) over a ConformerEncoder -- BASELINE config 4."""
    s = conformer_encoder_shapes(num_blocks=num_blocks, d=d)
    s["ctc.ctc_lo.weight"] = (vocab_size, d)
    s["ctc.ctc_lo.bias"] = (vocab_size,)
    return s


def count_params(shapes: Shapes) -> int:
    n = 0
    for shp in shapes.values():
        c = 1
        for d in shp:
            c *= d
        n += c
    return n
