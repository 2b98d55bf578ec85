// Package nanovllm — CudaModelRunner: the cgo shim that makes libnvllm_b200.so a drop-in
// ModelRunner for the unchanged Go engine.
//
// Drop this file into the reference's nanovllm/ package (next to tensor_model_runner.go,
// which it replaces at the call site) and build with
//
//	CGO_CFLAGS="-I<repo>/include" CGO_LDFLAGS="-L<repo>/nano-vllm-go_b200 -lnvllm_b200" go build ./...
//
// It implements exactly the interface of nanovllm/model_runner.go:9-16
//
//	Run(seqs []*Sequence, isPrefill bool) ([]int, error)
//	Close() error
//
// plus the extras of TensorModelRunner (tensor_model_runner.go:35-52,100-111) and
// GenericModelRunner (purego/generic_runner.go:184-196).  The image this library was built
// in has no Go toolchain, so this file is shipped UNCOMPILED (INTEGRATION.md says so);
// nano-vllm-go_b200/runner.py is the same shim over ctypes and is what the GPU tests exercise.
// tests/test_go_shim.py statically checks every selector used below (model.X, b.X.Y, mc.X,
// tensor.X, engineCfg.X, s.X, C.X) against the reference's Go declarations and include/nvllm.h.
package nanovllm

/*
#include <stdlib.h>
#include "nvllm.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"

	"nano-vllm-go/purego/tensor"
)

// CudaModelRunner implements ModelRunner on one B200.
type CudaModelRunner struct {
	h           *C.nvllm_handle
	cfg         *tensor.ModelConfig
	// sampling defaults of NewTensorModelRunner (tensor_model_runner.go:27-31 ->
	// tensor.DefaultSamplingParams, sampling.go:17-24): temperature 1.0, repetition penalty 1.2.
	// temperature <= 0 selects the greedy first-max rule of cmd/ask/main.go:389-402 (deviation:
	// the reference runner treats 0 as "no temperature scaling" and still draws).
	temperature float32

	// reusable flattening buffers (no Go pointer is retained by the library)
	tok, cuSeq, cached, bt, cuBt, out []C.int32_t
	temp                              []C.float
}

func lastError(h *C.nvllm_handle) error {
	return fmt.Errorf("nvllm: %s", C.GoString(C.nvllm_last_error(h)))
}

// fillConfig maps tensor.ModelConfig (purego/tensor/config.go:58-122) and the engine's pool
// sizing (nanovllm/config.go:9-20) onto nvllm_config.
func fillConfig(mc *tensor.ModelConfig, engineCfg *Config, device int, bf16 bool, tied bool) C.nvllm_config {
	var cc C.nvllm_config
	cc.abi_version = C.NVLLM_ABI_VERSION
	cc.vocab_size = C.int32_t(mc.VocabSize)
	cc.hidden = C.int32_t(mc.Hidden)
	cc.num_layers = C.int32_t(mc.NumLayers)
	cc.num_heads = C.int32_t(mc.NumHeads)
	cc.num_kv_heads = C.int32_t(mc.NumKVHeads)
	cc.head_dim = C.int32_t(mc.HeadDim)
	cc.ffn_dim = C.int32_t(mc.FFNDim)
	cc.max_seq_len = C.int32_t(mc.MaxSeqLen)
	cc.attention_type = C.int32_t(attnEnum(mc.AttentionType))
	cc.norm_type = C.int32_t(normEnum(mc.NormType))
	cc.position_type = C.int32_t(posEnum(mc.PositionType))
	cc.activation_type = C.int32_t(actEnum(mc.ActivationType))
	cc.block_style = C.int32_t(blockEnum(mc.BlockStyle))
	if tied {
		cc.tied_embedding = 1
	}
	if mc.UseMoE {
		cc.use_moe = 1
		cc.num_experts = C.int32_t(mc.NumExperts)
		cc.num_experts_per_tok = C.int32_t(mc.NumExpertsPerTok)
	}
	cc.norm_eps = C.float(mc.NormEps)
	cc.embedding_multiplier = C.float(mc.EmbeddingMultiplier)
	cc.attention_multiplier = C.float(mc.AttentionMultiplier)
	cc.residual_multiplier = C.float(mc.ResidualMultiplier)
	cc.logits_scaling = C.float(mc.LogitsScaling)
	cc.rope_base = C.double(mc.RoPEBase)
	if bf16 {
		cc.precision = C.NVLLM_PREC_BF16
	}
	cc.num_kv_blocks = C.int32_t(engineCfg.NumKVCacheBlocks)
	cc.max_num_seqs = C.int32_t(engineCfg.MaxNumSeqs)
	cc.max_batched_tokens = C.int32_t(engineCfg.MaxNumBatchedTokens)
	cc.tp_size = 1
	cc.device = C.int32_t(device)
	// Granite-4 hybrid (config.go:100-115); which layers are Mamba2 goes through
	// nvllm_set_layer_types (setLayerTypes below)
	cc.mamba_num_heads = C.int32_t(mc.Mamba2NumHeads)
	cc.mamba_head_dim = C.int32_t(mc.Mamba2HeadDim)
	cc.mamba_state_size = C.int32_t(mc.Mamba2StateSize)
	cc.mamba_n_groups = C.int32_t(mc.Mamba2NGroups)
	cc.mamba_conv_kernel = C.int32_t(mc.Mamba2ConvKernel)
	return cc
}

// setLayerTypes tells the library which blocks are Mamba2 (ModelConfig.HybridLayers,
// "mamba" / "mamba2" as in generic_model.go:74-75).  The library keeps the SSM state per KV
// block, so hybrid models need nothing else at the boundary — but the engine must not use the
// prefix cache or chunked prefill for them (the reference's causal conv restarts per call).
func setLayerTypes(h *C.nvllm_handle, mc *tensor.ModelConfig) error {
	if len(mc.HybridLayers) == 0 {
		return nil
	}
	types := make([]C.int32_t, len(mc.HybridLayers))
	for i, t := range mc.HybridLayers {
		if t == "mamba" || t == "mamba2" {
			types[i] = 1
		}
	}
	if rc := C.nvllm_set_layer_types(h, &types[0], C.int(len(types))); rc != 0 {
		return lastError(h)
	}
	return nil
}

// NewCudaModelRunnerFromDir is the constructor to prefer: only config.json goes through the
// reference's Go code (tensor.LoadModelConfig, generic_loader.go:808-973); the weights are read
// by the library itself, straight from model.safetensors or the sharded index into device
// memory (nvllm_load_safetensors), instead of being decoded to fp32 Go slices first
// (tensor.LoadModelFromDirectory, generic_loader.go:1016-1163).
func NewCudaModelRunnerFromDir(modelDir string, engineCfg *Config, device int, bf16 bool) (*CudaModelRunner, error) {
	mc, err := tensor.LoadModelConfig(modelDir + "/config.json")
	if err != nil {
		if mc, err = tensor.LoadModelConfig(modelDir + "/model_info.json"); err != nil {
			return nil, fmt.Errorf("no valid config found: %w", err)
		}
	}
	var arch C.int
	switch mc.Architecture {
	case tensor.ArchGPT2:
		arch = C.NVLLM_ARCH_GPT2
	case tensor.ArchFalcon:
		arch = C.NVLLM_ARCH_FALCON
	case tensor.ArchLlama:
		arch = C.NVLLM_ARCH_LLAMA
	case tensor.ArchGranite:
		if !mc.UseMoE {
			return nil, fmt.Errorf("granite hybrid (Mamba2) checkpoints are not supported by the CUDA runner")
		}
		arch = C.NVLLM_ARCH_GRANITE_MOE
	default:
		return nil, fmt.Errorf("unsupported architecture: %s", mc.Architecture)
	}
	// an untied config whose checkpoint has no lm_head falls back to the embedding inside the
	// library, like generic_loader.go:255-259
	cc := fillConfig(mc, engineCfg, device, bf16, mc.TiedEmbedding)
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var h *C.nvllm_handle
	if rc := C.nvllm_create(&cc, nil, &h); rc != 0 {
		return nil, lastError(nil)
	}
	fail := func() (*CudaModelRunner, error) {
		e := lastError(h)
		C.nvllm_destroy(h)
		return nil, e
	}
	if rc := C.nvllm_set_sampling(h, C.float(1.0), C.int(0), C.float(1.2)); rc != 0 {
		return fail()
	}
	dir := C.CString(modelDir)
	defer C.free(unsafe.Pointer(dir))
	if rc := C.nvllm_load_safetensors(h, dir, arch, nil, nil); rc != 0 {
		return fail()
	}
	if rc := C.nvllm_finalize(h); rc != 0 {
		return fail()
	}
	return &CudaModelRunner{h: h, cfg: mc, temperature: 1.0}, nil
}

// NewCudaModelRunner replaces NewTensorModelRunner (tensor_model_runner.go:21-32): it loads
// the model with the reference's own loader and hands every tensor, in the layout the loader
// leaves it in (generic_loader.go:184-265,353-604), to the library.
func NewCudaModelRunner(modelDir string, engineCfg *Config, device int, bf16 bool) (*CudaModelRunner, error) {
	model, err := tensor.LoadModelFromDirectory(modelDir)
	if err != nil {
		return nil, fmt.Errorf("failed to load model: %w", err)
	}
	mc := model.Config
	cc := fillConfig(mc, engineCfg, device, bf16, model.LMHead == nil || mc.TiedEmbedding)

	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var h *C.nvllm_handle
	if rc := C.nvllm_create(&cc, nil, &h); rc != 0 {
		return nil, lastError(nil)
	}
	r := &CudaModelRunner{h: h, cfg: mc, temperature: 1.0}
	if rc := C.nvllm_set_sampling(h, C.float(1.0), C.int(0), C.float(1.2)); rc != 0 {
		e := lastError(h)
		C.nvllm_destroy(h)
		return nil, e
	}
	if e := setLayerTypes(h, mc); e != nil {
		C.nvllm_destroy(h)
		return nil, e
	}
	set := func(id C.int, layer int, t *tensor.Tensor) error {
		if t == nil || len(t.Data) == 0 {
			return nil // optional tensors (biases) are simply not sent
		}
		if rc := C.nvllm_set_weight(h, id, C.int(layer), C.NVLLM_DT_F32,
			unsafe.Pointer(&t.Data[0]), C.uint64_t(len(t.Data))); rc != 0 {
			return lastError(h)
		}
		return nil
	}
	steps := []error{
		set(C.NVLLM_W_TOKEN_EMBEDDING, -1, model.TokenEmbedding),
		set(C.NVLLM_W_POS_EMBEDDING, -1, model.PosEmbedding),
		set(C.NVLLM_W_FINAL_NORM_W, -1, model.LNFinal.Weight),
		set(C.NVLLM_W_FINAL_NORM_B, -1, model.LNFinal.Bias),
	}
	if cc.tied_embedding == 0 {
		steps = append(steps, set(C.NVLLM_W_LM_HEAD, -1, model.LMHead))
	}
	for l, b := range model.Blocks {
		steps = append(steps, r.setBlock(set, l, b)...)
	}
	for _, e := range steps {
		if e != nil {
			C.nvllm_destroy(h)
			return nil, e
		}
	}
	if rc := C.nvllm_finalize(h); rc != 0 {
		e := lastError(h)
		C.nvllm_destroy(h)
		return nil, e
	}
	return r, nil
}

// setBlock sends one GenericBlock (generic_model.go:22-31).
func (r *CudaModelRunner) setBlock(set func(C.int, int, *tensor.Tensor) error, l int, b *tensor.GenericBlock) []error {
	var es []error
	if b.AttnLN != nil {
		es = append(es, set(C.NVLLM_W_ATTN_NORM_W, l, b.AttnLN.Weight), set(C.NVLLM_W_ATTN_NORM_B, l, b.AttnLN.Bias))
	}
	if b.InputLN != nil { // parallel blocks (Falcon)
		es = append(es, set(C.NVLLM_W_ATTN_NORM_W, l, b.InputLN.Weight), set(C.NVLLM_W_ATTN_NORM_B, l, b.InputLN.Bias))
	}
	if b.FFNLN != nil {
		es = append(es, set(C.NVLLM_W_FFN_NORM_W, l, b.FFNLN.Weight), set(C.NVLLM_W_FFN_NORM_B, l, b.FFNLN.Bias))
	}
	switch a := b.Attention.(type) {
	case *tensor.MultiHeadAttention:
		es = append(es, set(C.NVLLM_W_Q, l, a.QWeight), set(C.NVLLM_W_K, l, a.KWeight), set(C.NVLLM_W_V, l, a.VWeight),
			set(C.NVLLM_W_O, l, a.OutWeight), set(C.NVLLM_W_Q_BIAS, l, a.QBias), set(C.NVLLM_W_K_BIAS, l, a.KBias),
			set(C.NVLLM_W_V_BIAS, l, a.VBias), set(C.NVLLM_W_O_BIAS, l, a.OutBias))
	case *tensor.GroupedQueryAttention:
		es = append(es, set(C.NVLLM_W_Q, l, a.QWeight), set(C.NVLLM_W_K, l, a.KWeight), set(C.NVLLM_W_V, l, a.VWeight),
			set(C.NVLLM_W_O, l, a.OutWeight))
	case *tensor.MultiQueryAttention:
		es = append(es, set(C.NVLLM_W_Q, l, a.QWeight), set(C.NVLLM_W_KV, l, a.KVWeight), set(C.NVLLM_W_O, l, a.OutWeight),
			set(C.NVLLM_W_Q_BIAS, l, a.QBias), set(C.NVLLM_W_O_BIAS, l, a.OutBias))
		if a.KVBias != nil { // [K | V] halves (mqa.go:20,120-125): the library keeps K and V biases apart
			hd := a.HeadDim
			es = append(es,
				set(C.NVLLM_W_K_BIAS, l, &tensor.Tensor{Data: a.KVBias.Data[:hd], Shape: []int{hd}}),
				set(C.NVLLM_W_V_BIAS, l, &tensor.Tensor{Data: a.KVBias.Data[hd : 2*hd], Shape: []int{hd}}))
		}
	case *tensor.Mamba2Layer:
		// kept as loaded, [out, in] (mamba2.go:87,178 multiply by the transpose); XProj / DtProj are
		// never read by Forward and ConvCache / SSMState are run-time state: not sent
		es = append(es, set(C.NVLLM_W_MAMBA_IN_PROJ, l, a.InProj), set(C.NVLLM_W_MAMBA_CONV_W, l, a.ConvWeight),
			set(C.NVLLM_W_MAMBA_CONV_B, l, a.ConvBias), set(C.NVLLM_W_MAMBA_DT_BIAS, l, a.DeltaBias),
			set(C.NVLLM_W_MAMBA_A_LOG, l, a.ALog), set(C.NVLLM_W_MAMBA_D, l, a.D),
			set(C.NVLLM_W_MAMBA_NORM, l, a.Norm), set(C.NVLLM_W_MAMBA_OUT_PROJ, l, a.OutProj))
	}
	if b.MoE != nil {
		es = append(es, set(C.NVLLM_W_MOE_ROUTER, l, b.MoE.Router), set(C.NVLLM_W_MOE_INPUT, l, b.MoE.InputLinear),
			set(C.NVLLM_W_MOE_OUTPUT, l, b.MoE.OutputLinear))
	} else if b.FFN != nil {
		es = append(es, set(C.NVLLM_W_FFN_W1, l, b.FFN.W1), set(C.NVLLM_W_FFN_W2, l, b.FFN.W2),
			set(C.NVLLM_W_FFN_B2, l, b.FFN.B2))
		if r.cfg.ActivationType != tensor.ActivationSwiGLU {
			es = append(es, set(C.NVLLM_W_FFN_B1, l, b.FFN.B1))
		}
	}
	return es
}

// Run implements ModelRunner.Run (model_runner.go:11): one next token per scheduled
// sequence, same order.  The whole batch is one device step.
func (r *CudaModelRunner) Run(seqs []*Sequence, isPrefill bool) ([]int, error) {
	n := len(seqs)
	if n == 0 {
		return nil, nil
	}
	r.tok, r.bt = r.tok[:0], r.bt[:0]
	r.cuSeq = append(r.cuSeq[:0], 0)
	r.cuBt = append(r.cuBt[:0], 0)
	r.cached, r.temp = r.cached[:0], r.temp[:0]
	for _, s := range seqs {
		for _, t := range s.TokenIDs {
			r.tok = append(r.tok, C.int32_t(t))
		}
		r.cuSeq = append(r.cuSeq, C.int32_t(len(r.tok)))
		for _, b := range s.BlockTable {
			r.bt = append(r.bt, C.int32_t(b))
		}
		r.cuBt = append(r.cuBt, C.int32_t(len(r.bt)))
		r.cached = append(r.cached, C.int32_t(s.NumCachedTokens))
		r.temp = append(r.temp, C.float(r.temperature))
	}
	if cap(r.out) < n {
		r.out = make([]C.int32_t, n)
	}
	r.out = r.out[:n]
	prefill := C.int(0)
	if isPrefill {
		prefill = 1
	}
	// cgo may move this goroutine between OS threads between calls; every entry point of the
	// library does cudaSetDevice itself, so no thread pinning is needed here.
	rc := C.nvllm_run(r.h, C.int(n), prefill, i32ptr(r.tok), i32ptr(r.cuSeq), i32ptr(r.cached), i32ptr(r.bt),
		i32ptr(r.cuBt), &r.temp[0], nil, &r.out[0], nil, nil)
	if rc != 0 {
		return nil, lastError(r.h)
	}
	res := make([]int, n)
	for i := range res {
		res[i] = int(r.out[i])
	}
	return res, nil
}

// i32ptr is &s[0] that tolerates an empty slice (a sequence without blocks or tokens is
// rejected by the library with NVLLM_ERR_INVALID instead of a Go panic here).
func i32ptr(s []C.int32_t) *C.int32_t {
	if len(s) == 0 {
		return nil
	}
	return &s[0]
}

// Close implements ModelRunner.Close (model_runner.go:15).
func (r *CudaModelRunner) Close() error {
	if r.h != nil {
		C.nvllm_destroy(r.h)
		r.h = nil
	}
	return nil
}

// SetSamplingParams mirrors tensor_model_runner.go:35-42 (repetition penalty 1.2 by default).
func (r *CudaModelRunner) SetSamplingParams(temperature, topP float32, topK int) {
	r.SetSamplingParamsWithRepetition(temperature, topP, topK, 1.2)
}

// SetSamplingParamsWithRepetition mirrors tensor_model_runner.go:45-52.
func (r *CudaModelRunner) SetSamplingParamsWithRepetition(temperature, topP float32, topK int, rep float32) {
	r.temperature = temperature
	C.nvllm_set_sampling(r.h, C.float(topP), C.int(topK), C.float(rep))
}

// SetSpeculation toggles the library's speculative next decode step (include/nvllm.h:
// nvllm_set_speculation; on by default).  Not part of the reference's method set.
func (r *CudaModelRunner) SetSpeculation(on bool) {
	v := 0
	if on {
		v = 1
	}
	C.nvllm_set_speculation(r.h, C.int(v))
}

// ClearCache / ClearAllCaches (tensor_model_runner.go:100-111): KV state lives in the paged
// pool and is addressed only through BlockTable, so there is nothing to free per sequence.
func (r *CudaModelRunner) ClearCache(seqID int64) {}
func (r *CudaModelRunner) ClearAllCaches()        {}

// GetVocabSize / GetModelConfig (purego/generic_runner.go:184-196).
func (r *CudaModelRunner) GetVocabSize() int                  { return r.cfg.VocabSize }
func (r *CudaModelRunner) GetModelConfig() *tensor.ModelConfig { return r.cfg }

func attnEnum(t tensor.AttentionType) int {
	switch t {
	case tensor.AttentionGQA:
		return C.NVLLM_ATTN_GQA
	case tensor.AttentionMQA:
		return C.NVLLM_ATTN_MQA
	}
	return C.NVLLM_ATTN_MHA
}
func normEnum(t tensor.NormType) int {
	if t == tensor.NormRMS {
		return C.NVLLM_NORM_RMS
	}
	return C.NVLLM_NORM_LAYER
}
func posEnum(t tensor.PositionType) int {
	switch t {
	case tensor.PositionRoPE:
		return C.NVLLM_POS_ROPE
	case tensor.PositionLearned:
		return C.NVLLM_POS_LEARNED
	}
	return C.NVLLM_POS_NONE
}
func actEnum(t tensor.ActivationType) int {
	if t == tensor.ActivationSwiGLU {
		return C.NVLLM_ACT_SWIGLU
	}
	return C.NVLLM_ACT_GELU
}
func blockEnum(t tensor.BlockStyle) int {
	if t == tensor.BlockParallel {
		return C.NVLLM_BLOCK_PARALLEL
	}
	return C.NVLLM_BLOCK_SEQUENTIAL
}
