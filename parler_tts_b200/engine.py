"""Host-side driver of the CUDA decode path: packs reference-layout weights into kernel-friendly device buffers,
owns the static KV caches / token matrix / control block, and enqueues the C-ABI calls (optionally as a CUDA graph).

PyTorch is used for device memory, streams and graphs only; all arithmetic on the hot path happens in
libparler_b200.so.  Reference behaviour being replaced: /root/reference/parler_tts/modeling_parler_tts.py
  ParlerTTSDecoder.forward 794-978, ParlerTTSForCausalLM.forward 1096-1185, prepare_inputs_for_generation 2081-2148,
  and the transformers greedy_search / sample loop (call sites 2582-2620).
"""
import ctypes as C
from dataclasses import dataclass

import torch

from . import cabi


@dataclass
class DecoderDims:
    vocab_size: int = 1088
    hidden_size: int = 1024
    num_hidden_layers: int = 24
    num_attention_heads: int = 16
    ffn_dim: int = 4096
    num_codebooks: int = 9
    max_position_embeddings: int = 4096
    pad_token_id: int = 1024
    eos_token_id: int = 1024
    bos_token_id: int = 1025
    activation_function: str = "gelu"

    @classmethod
    def from_config(cls, cfg):
        return cls(**{f: getattr(cfg, f) for f in cls.__dataclass_fields__})

    @property
    def head_dim(self):
        return self.hidden_size // self.num_attention_heads


class DecoderWeights:
    """Packed, contiguous device buffers built once from a state-dict with the reference's key names
    (`model.decoder.layers.{i}.self_attn.q_proj.weight`, ..., `lm_heads.{k}.weight`)."""

    def __init__(self, sd, dims: DecoderDims, device, dtype):
        self.dims, self.device, self.dtype = dims, torch.device(device), dtype
        L, H, K = dims.num_hidden_layers, dims.hidden_size, dims.num_codebooks

        def mat(*names):  # stacked [sum out][in] matrix in the compute dtype
            return torch.cat([sd[n].detach() for n in names], dim=0).to(self.device, dtype).contiguous()

        def vec(name):
            return sd[name].detach().to(self.device, torch.float32).contiguous()

        self.embed = torch.stack([sd[f"model.decoder.embed_tokens.{k}.weight"].detach() for k in range(K)]).to(
            self.device, torch.float32).contiguous()
        assert self.embed.shape == (K, dims.vocab_size + 1, H), self.embed.shape
        self.pos = vec("model.decoder.embed_positions.weights")
        self.lnf_g, self.lnf_b = vec("model.decoder.layer_norm.weight"), vec("model.decoder.layer_norm.bias")
        self.w_heads = mat(*[f"lm_heads.{k}.weight" for k in range(K)])
        self.layers = []
        arr = (cabi.PttsLayer * L)()
        for i in range(L):
            p = f"model.decoder.layers.{i}."
            t = dict(
                w_qkv=mat(p + "self_attn.q_proj.weight", p + "self_attn.k_proj.weight", p + "self_attn.v_proj.weight"),
                w_o=mat(p + "self_attn.out_proj.weight"),
                w_qc=mat(p + "encoder_attn.q_proj.weight"),
                w_kvc=mat(p + "encoder_attn.k_proj.weight", p + "encoder_attn.v_proj.weight"),
                w_oc=mat(p + "encoder_attn.out_proj.weight"),
                w_fc1=mat(p + "fc1.weight"),
                w_fc2=mat(p + "fc2.weight"),
                ln1_g=vec(p + "self_attn_layer_norm.weight"), ln1_b=vec(p + "self_attn_layer_norm.bias"),
                ln2_g=vec(p + "encoder_attn_layer_norm.weight"), ln2_b=vec(p + "encoder_attn_layer_norm.bias"),
                ln3_g=vec(p + "final_layer_norm.weight"), ln3_b=vec(p + "final_layer_norm.bias"),
            )
            self.layers.append(t)
            for name, tensor in t.items():
                setattr(arr[i], name, tensor.data_ptr())
        self._layer_array = arr
        m = cabi.PttsModel()
        m.n_layers, m.hidden, m.n_heads, m.head_dim = L, H, dims.num_attention_heads, dims.head_dim
        m.ffn, m.n_codebooks, m.vocab, m.max_pos = dims.ffn_dim, K, dims.vocab_size, self.pos.shape[0]
        m.wdtype, m.act_fn = cabi.dtype_enum(dtype), cabi.ACT[dims.activation_function]
        m.pad_id, m.eos_id, m.bos_id = dims.pad_token_id, dims.eos_token_id, dims.bos_token_id
        m.embed, m.pos = self.embed.data_ptr(), self.pos.data_ptr()
        m.lnf_g, m.lnf_b, m.w_heads = self.lnf_g.data_ptr(), self.lnf_b.data_ptr(), self.w_heads.data_ptr()
        m.layers = C.cast(arr, C.POINTER(cabi.PttsLayer))
        self.c_model = m

    def weight_elements(self):
        """Matrix + LayerNorm elements streamed by one decode step (W)."""
        d = self.dims
        H, F = d.hidden_size, d.ffn_dim
        return d.num_hidden_layers * (6 * H * H + 2 * H * F + 6 * H) + 2 * H + d.num_codebooks * d.vocab_size * H


class GenState:
    """All device state of one batch of utterances: static KV caches, token matrix, delay mask, control block."""

    def __init__(self, weights: DecoderWeights, B, S, P, T0, max_len, impl=cabi.IMPL_FAST, do_sample=False, top_k=50,
                 temperature=1.0, suppress_from=-1, guidance_scale=1.0, seed=0, row_offset=0, enc_mask=None,
                 prompt_mask=None, kv_dtype=None, min_length=0, top_p=1.0, repetition_penalty=1.0, pad_token_id=None,
                 eos_token_id=None, bos_token_id=None, self_attn_kernel=0):
        d, dev, dt = weights.dims, weights.device, weights.dtype
        self.weights = weights
        self.B, self.S, self.P, self.T0, self.max_len = B, S, P, T0, max_len
        self.cfg = guidance_scale is not None and guidance_scale > 1
        self.Bu = B // 2 if self.cfg else B
        K, H, F, V = d.num_codebooks, d.hidden_size, d.ffn_dim, d.vocab_size
        heads, hd, L = d.num_attention_heads, d.head_dim, d.num_hidden_layers
        self.c_max = P + max_len
        rows = B * max(P + T0, S, 1)
        z = lambda *shape, dtype=dt: torch.zeros(*shape, device=dev, dtype=dtype)  # noqa: E731
        self.ctrl = z(8, dtype=torch.int32)
        # self-attention cache: the weight dtype, or (kv_dtype="fp8") e4m3 bytes + per-(layer, K|V, head) scales + a
        # staging cache in the weight dtype that only the prefill touches
        self.kv_dtype = "fp8" if kv_dtype in ("fp8", "fp8_e4m3", torch.float8_e4m3fn) else None
        if self.kv_dtype == "fp8":
            if dt != torch.bfloat16 or impl != cabi.IMPL_FAST or hd != 64:
                raise cabi.ParlerB200Error("kv_dtype='fp8' needs bf16 weights, the fast kernel family and head_dim 64")
            self.kv_self = z(L, 2, B, heads, self.c_max, hd, dtype=torch.uint8)
            self.kv_scale = torch.ones(L, 2, heads, device=dev, dtype=torch.float32)
            self.kv_stage = z(L, 2, B, heads, P + T0, hd)
        else:
            self.kv_self = z(L, 2, B, heads, self.c_max, hd)
            self.kv_scale = self.kv_stage = None
        self.kv_cross = z(L, 2, B, heads, max(S, 1), hd)
        self.h = z(rows, H, dtype=torch.float32)
        self.xn, self.q, self.attn, self.ffn = z(rows, H), z(rows, H), z(rows, H), z(rows, F)
        self.logits = z(B * K, V, dtype=torch.float32)
        n_scratch = int(cabi.load().ptts_scratch_floats(C.byref(weights.c_model), rows))
        self.scratch = z(n_scratch, dtype=torch.float32)
        self.ids = z(self.Bu * K, max_len, dtype=torch.int64)
        self.mask = z(self.Bu * K, max_len, dtype=torch.int64)
        self.unfinished = torch.ones(self.Bu * K, device=dev, dtype=torch.int32)
        self.first_start = z(1, dtype=torch.int32)
        self.enc_mask = None if enc_mask is None else enc_mask.to(dev, torch.int32).contiguous()
        self.prompt_mask = None if prompt_mask is None else prompt_mask.to(dev, torch.int32).contiguous()
        s = cabi.PttsState()
        s.B, s.S, s.P, s.T0, s.c_max, s.max_len, s.impl = B, S, P, T0, self.c_max, max_len, impl
        s.do_sample, s.top_k, s.temperature = int(bool(do_sample)), int(top_k or 0), float(temperature)
        s.suppress_from = int(suppress_from if suppress_from is not None else -1)
        s.guidance_scale = float(guidance_scale) if self.cfg else 1.0
        s.seed, s.row_offset = int(seed) & 0xFFFFFFFFFFFFFFFF, int(row_offset)
        s.self_attn_kernel = int(self_attn_kernel or 0)
        s.kv_dtype = cabi.PTTS_FP8 if self.kv_dtype == "fp8" else cabi.dtype_enum(dt)
        s.kv_scale = None if self.kv_scale is None else self.kv_scale.data_ptr()
        s.kv_stage = None if self.kv_stage is None else self.kv_stage.data_ptr()
        s.min_length, s.top_p = int(min_length or 0), float(1.0 if top_p is None else top_p)
        s.repetition_penalty = float(1.0 if repetition_penalty is None else repetition_penalty)
        # generation-config ids (reference: generation_config.{pad,eos}_token_id drive the loop, 2586-2587; the
        # delay mask uses generation_config.{decoder_start,pad}_token_id, 2530-2535), defaulting to the decoder config's
        self.pad_id = d.pad_token_id if pad_token_id is None else int(pad_token_id)
        self.eos_id = d.eos_token_id if eos_token_id is None else int(eos_token_id)
        self.bos_id = d.bos_token_id if bos_token_id is None else int(bos_token_id)
        s.pad_id, s.eos_id = self.pad_id, self.eos_id
        for name in ("ctrl", "kv_self", "kv_cross", "h", "xn", "q", "attn", "ffn", "logits", "scratch", "ids", "mask",
                     "unfinished"):
            setattr(s, name, getattr(self, name).data_ptr())
        s.enc_mask = None if self.enc_mask is None else self.enc_mask.data_ptr()
        s.prompt_mask = None if self.prompt_mask is None else self.prompt_mask.data_ptr()
        self.c_state = s
        self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self.graph, self.graph_launches, self.replays = None, 0, 0
        self.graph_multi, self.steps_per_graph = None, 1  # optional second graph holding several consecutive steps
        self._ctrl_host = torch.zeros(8, dtype=torch.int32).pin_memory() if dev.type == "cuda" else None

    # ---- control block helpers
    def reset(self, start_ids, seed=None):
        """start_ids (Bu*K, T0) int64: builds the delay mask and initialises ids / control block (+ Philox key)."""
        if seed is not None:
            self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        d = self.weights.dims
        K = d.num_codebooks
        start = start_ids.to(self.ids.device, torch.int64).contiguous()
        assert start.shape == (self.Bu * K, self.T0), (start.shape, self.Bu, K, self.T0)
        cabi.call("ptts_build_delay_mask", cabi.ptr(start), self.Bu, K, self.T0, self.max_len, self.bos_id,
                  self.pad_id, cabi.ptr(self.mask), cabi.ptr(self.first_start), cabi.stream_ptr())
        self.ids.zero_()
        self.ids[:, : self.T0] = start
        self.unfinished.fill_(1)
        lo, hi = self.seed & 0xFFFFFFFF, self.seed >> 32
        to_i32 = lambda v: v - (1 << 32) if v >= (1 << 31) else v  # noqa: E731
        ctrl = torch.tensor([self.T0, 0, 0, self.Bu * K, 0, 0, to_i32(lo), to_i32(hi)], dtype=torch.int32)
        self.ctrl.copy_(ctrl.to(self.ctrl.device))

    def read_ctrl(self):
        c = self.ctrl.cpu().tolist()
        return dict(len=c[0], done=c[1], stop_len=c[2], n_unfinished=c[3], steps=c[5])


class DecoderEngine:
    steps_per_graph = 1  # decode steps per CUDA-graph replay of run_loop (see capture_multi)

    def __init__(self, weights: DecoderWeights):
        self.weights = weights
        self.lib = cabi.load()

    def _m(self):
        return C.byref(self.weights.c_model)

    def cross_kv(self, st: GenState, enc):
        """enc (B, S, H) projected + masked encoder states, compute dtype."""
        enc = enc.to(self.weights.device, self.weights.dtype).contiguous()
        cabi.call("ptts_cross_kv", self._m(), C.byref(st.c_state), cabi.ptr(enc), cabi.stream_ptr())

    def prefill(self, st: GenState, prompt_table=None, prompt_ids=None, prompt_hidden=None, all_logits=False,
                run_epilogue=True):
        d = self.weights.dims
        out = None
        if all_logits:
            out = torch.empty(st.B * d.num_codebooks, st.P + st.T0, d.vocab_size, device=self.weights.device,
                              dtype=torch.float32)
        keep = [t for t in (prompt_table, prompt_ids, prompt_hidden) if t is not None]
        cabi.call("ptts_prefill", self._m(), C.byref(st.c_state), cabi.ptr(prompt_table), cabi.ptr(prompt_ids),
                  cabi.ptr(prompt_hidden), cabi.ptr(out), int(run_epilogue), cabi.stream_ptr())
        del keep
        return out

    def step(self, st: GenState, run_epilogue=True):
        cabi.call("ptts_decode_step", self._m(), C.byref(st.c_state), int(run_epilogue), cabi.stream_ptr())

    def epilogue(self, st: GenState):
        cabi.call("ptts_step_epilogue", self._m(), C.byref(st.c_state), cabi.stream_ptr())

    # stream priority the step graph is captured with (kernel nodes keep the priority of the capturing stream): -1 puts
    # the loop's kernels in front of the chunked codec decode that shares the GPU with them (streaming.py); 0 = default
    graph_priority = 0

    def _capture_stream(self):
        return torch.cuda.Stream(priority=self.graph_priority) if self.graph_priority else None

    def capture(self, st: GenState):
        """Captures one whole decode step (all layers + heads + epilogue) into a CUDA graph.  Sequence length lives in
        the device control block, so the same graph is replayed for every step."""
        if st.graph is not None:
            return st.graph
        g = torch.cuda.CUDAGraph()
        n0 = self.lib.ptts_launch_count()
        with torch.cuda.graph(g, stream=self._capture_stream()):
            self.step(st)
        st.graph_launches = int(self.lib.ptts_launch_count() - n0)
        st.graph = g
        return g

    def capture_multi(self, st: GenState, steps):
        """A second graph of `steps` consecutive decode steps (the position counter lives on the device, so the steps
        of one replay advance like separate replays): removes the graph-to-graph hand-over between them.  Within the
        graph every kernel still waits (griddepcontrol.wait) for the one before it, the epilogue never releases its
        dependent early, so no kernel of step i + 1 can read the counter before step i's epilogue has written it."""
        if st.graph_multi is not None and st.steps_per_graph == steps:
            return st.graph_multi
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=self._capture_stream()):
            for _ in range(steps):
                self.step(st)
        st.graph_multi, st.steps_per_graph = g, steps
        return g

    def run_loop(self, st: GenState, use_graph=True, poll_every=16, max_steps=None, on_progress=None,
                 steps_per_graph=None):
        """Runs decode steps until the control block says done (or max_steps).  Returns the control block.
        on_progress(steps): called on the host after every batch of enqueued steps (streaming.ChunkPipeline hangs the
        overlapped chunk decode here); it must not synchronise with the loop's stream."""
        budget = st.max_len - st.T0 - 1 if max_steps is None else max_steps
        if budget <= 0:
            return st.read_ctrl()
        done_steps = 0
        g = None
        if use_graph:
            if st.graph is None:  # one eager step first: loads every kernel before stream capture starts
                self.step(st)
                done_steps = 1
            g = self.capture(st)
        U = int(steps_per_graph if steps_per_graph is not None else self.steps_per_graph)
        gm = self.capture_multi(st, U) if (g is not None and U > 1) else None
        ev = None
        while done_steps < budget:
            n = min(poll_every, budget - done_steps)
            left = n
            while gm is not None and left >= U:  # whole multi-step graphs first, single steps for the remainder
                gm.replay()
                st.replays += U
                left -= U
            for _ in range(left):
                if g is not None:
                    g.replay()
                    st.replays += 1
                else:
                    self.step(st)
            done_steps += n
            if on_progress is not None:
                on_progress(done_steps)
            if ev is not None:
                ev.synchronize()
                if int(st._ctrl_host[1]) != 0:
                    break
            st._ctrl_host.copy_(st.ctrl, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        return st.read_ctrl()
