//! `crates/core/src/quantization/b200_w4a16.rs` -- the shim a maintainer of vasia123/vllm-rust adds to route the W4A16
//! hot path (AWQ / AWQ-Marlin / GPTQ int4 linears) through `libw4a16_b200.so` on sm_100a.
//!
//! Shipped as source: this repository's image has no Rust toolchain, so the file is not compiled here.  What IS checked
//! here (tests/test_integration_rs.py): the `extern "C"` block (`b200_w4a16_ffi.rs`, generated from
//! `include/w4a16_b200.h`) matches the header symbol for symbol and type for type, and every `ffi::` symbol this file
//! calls exists in it.  The Python mirror (`vllm-rust_b200/quantization.py`) exercises the same calls on the GPU.
//!
//! Drop-in points (reference file:line):
//!   * `AwqWeightLoader::load_linear` (weight_loader.rs:686-702) returns `B200W4A16Linear` where it builds
//!     `AwqMarlinLinear::new` today; `GptqWeightLoader::load_linear` (weight_loader.rs:264-316) likewise.
//!   * `QuantizedLinear::forward` (config.rs:147-190) -> `w4a16_forward_into`; pooled outputs as in
//!     `marlin_gemm_pooled` (marlin_cuda.rs:1655-1753).
//!   * `RowParallelLinear::forward` (parallel_layers.rs:275-313) -> `forward_row`.
//!   * the decode layer loop (qwen3_quantized.rs:704-709, 532-551) under CUDA-graph replay (engine/cuda_graph.rs)
//!     -> `DecodeChain` (one persistent launch for a fixed sequence of linears, M <= 4).
#[path = "b200_w4a16_ffi.rs"]
mod ffi;

use std::collections::HashMap;
use std::ffi::CStr;
use std::os::raw::{c_int, c_void};

use candle_core::cuda_backend::cudarc::driver::DevicePtr;
use candle_core::cuda_backend::{CudaStorageSlice, WrapErr};
use candle_core::{DType, Device, Result, Storage, Tensor};

use super::config::QuantizedLinear;

pub const FORMAT_AWQ: c_int = 0;
pub const FORMAT_GPTQ_ZP: c_int = 1;
pub const FORMAT_GPTQ_SYM: c_int = 2;
pub const TP_NONE: c_int = 0;
pub const TP_COLUMN: c_int = 1;
pub const TP_ROW: c_int = 2;
pub const EPILOGUE_SWIGLU: c_int = 1;

/// Every failure carries the library's message, like the launchers' `bail!` (marlin_cuda.rs:546-564).
fn check(rc: c_int) -> Result<()> {
    if rc == 0 {
        return Ok(());
    }
    let msg = unsafe { CStr::from_ptr(ffi::w4a16_last_error()) }.to_string_lossy().into_owned();
    candle_core::bail!("{msg}")
}

fn dtype_code(d: DType) -> Result<c_int> {
    match d {
        DType::F16 => Ok(0),
        DType::BF16 => Ok(1),
        DType::F32 => Ok(2),
        other => candle_core::bail!("b200_w4a16: unsupported scales / bias dtype {other:?}"),
    }
}

/// Raw device pointer of a contiguous CUDA tensor (the pattern of nccl.rs:1233-1265: the guard of `device_ptr` is
/// dropped at once; ordering is by the stream every call below is enqueued on).
fn dev_ptr(t: &Tensor) -> Result<*const c_void> {
    if !t.is_contiguous() {
        candle_core::bail!("b200_w4a16: tensor must be contiguous");
    }
    let (storage, layout) = t.storage_and_layout();
    let cs = match &*storage {
        Storage::Cuda(cs) => cs,
        _ => candle_core::bail!("b200_w4a16: tensor must be on CUDA (there is no CPU fallback)"),
    };
    let off = layout.start_offset();
    macro_rules! ptr_of {
        ($s:expr) => {{
            let view = $s.slice(off..);
            let (p, _guard) = view.device_ptr($s.stream());
            p as usize as *const c_void
        }};
    }
    Ok(match &cs.slice {
        CudaStorageSlice::U8(s) => ptr_of!(s),
        CudaStorageSlice::U32(s) => ptr_of!(s),
        CudaStorageSlice::I64(s) => ptr_of!(s),
        CudaStorageSlice::BF16(s) => ptr_of!(s),
        CudaStorageSlice::F16(s) => ptr_of!(s),
        CudaStorageSlice::F32(s) => ptr_of!(s),
        _ => candle_core::bail!("b200_w4a16: unsupported storage dtype"),
    })
}

/// The stream candle enqueues this device's kernels on (qwen3_quantized.rs:107-109 uses the same accessor).
fn stream_ptr(device: &Device) -> Result<*mut c_void> {
    match device {
        Device::Cuda(cd) => Ok(cd.cuda_stream().cu_stream() as *mut c_void),
        _ => candle_core::bail!("b200_w4a16: needs a CUDA device (there is no CPU fallback)"),
    }
}

fn gpu_ordinal(device: &Device) -> Result<c_int> {
    match device.location() {
        candle_core::DeviceLocation::Cuda { gpu_id } => Ok(gpu_id as c_int),
        _ => candle_core::bail!("b200_w4a16: needs a CUDA device (there is no CPU fallback)"),
    }
}

/// Drop-in for `AwqMarlinLinear` (awq_marlin.rs:431-667), `MarlinLinear` (marlin.rs:586-778) and `GptqLinear`
/// (gptq.rs:254-471).
pub struct B200W4A16Linear {
    h: *mut ffi::W4a16Linear,
    format: c_int,
    has_bias: bool,
    device: Device,
}
// The weights are read-only after load_weights; the handle's scratch is per handle and the engine drives one forward at
// a time per handle (the same contract the reference's MarlinLinear has through its workspace tensor).
unsafe impl Send for B200W4A16Linear {}
unsafe impl Sync for B200W4A16Linear {}

impl B200W4A16Linear {
    /// in_features / out_features are the FULL (unsharded) sizes; `tp` = (mode, rank, size).
    pub fn new(in_features: usize, out_features: usize, group_size: i32, format: c_int, has_bias: bool, device: &Device,
               tp: (c_int, c_int, c_int)) -> Result<Self> {
        let mut h = std::ptr::null_mut();
        check(unsafe {
            ffi::w4a16_create_shard(&mut h, in_features as i64, out_features as i64, group_size, format, has_bias as c_int,
                                    gpu_ordinal(device)?, tp.0, tp.1, tp.2)
        })?;
        Ok(Self { h, format, has_bias, device: device.clone() })
    }

    pub fn raw(&self) -> *mut ffi::W4a16Linear {
        self.h
    }

    /// gate_proj + up_proj + fused_swiglu (qwen3_quantized.rs:501-551) as one handle.
    pub fn set_swiglu(&mut self) -> Result<()> {
        check(unsafe { ffi::w4a16_set_epilogue(self.h, EPILOGUE_SWIGLU) })
    }

    /// One sibling projection (q / k / v, gate / up) into output columns [n_begin, n_begin + its width).
    pub fn load_part(&mut self, n_begin: usize, w: &HashMap<String, Tensor>) -> Result<()> {
        self.load_impl(w, Some(n_begin))
    }

    fn load_impl(&mut self, w: &HashMap<String, Tensor>, part: Option<usize>) -> Result<()> {
        // keys and checkpoint layouts exactly as awq_marlin.rs:592-664 / gptq.rs:435-454: the repack happens in the library
        let get = |k: &str| w.get(k).ok_or_else(|| candle_core::Error::Msg(format!("Missing '{k}'")));
        let qweight = get("qweight")?;
        let scales = get("scales")?;
        let stream = stream_ptr(&self.device)?;
        let (bias_p, bias_d) = match w.get("bias") {
            Some(b) if self.has_bias => (dev_ptr(b)?, dtype_code(b.dtype())?),
            _ => (std::ptr::null(), 1),
        };
        let qzeros_p = match w.get("qzeros") {
            Some(z) => dev_ptr(z)? as *const u32,
            None if self.format == FORMAT_GPTQ_SYM => std::ptr::null(),
            None => candle_core::bail!("Missing 'qzeros'"),
        };
        // g_idx: device int32 [K]; the converted tensor must outlive the call (the library reads it on the host side
        // of the call and synchronises the stream once, w4a16_b200.h)
        let g_idx_t = match w.get("g_idx") {
            Some(g) => Some(g.to_dtype(DType::U32)?.contiguous()?),
            None => None,
        };
        let g_idx_p = match &g_idx_t {
            Some(g) => dev_ptr(g)? as *const i32,
            None => std::ptr::null(),
        };
        let (qw, sc, sd) = (dev_ptr(qweight)? as *const u32, dev_ptr(scales)?, dtype_code(scales.dtype())?);
        let n_part = scales.dims()[1] as i64;
        check(unsafe {
            match (self.format == FORMAT_AWQ, part) {
                (true, None) => ffi::w4a16_load_awq(self.h, qw, qzeros_p, sc, sd, bias_p, bias_d, stream),
                (true, Some(n0)) => ffi::w4a16_load_awq_part(self.h, n0 as i64, n_part, qw, qzeros_p, sc, sd, bias_p, bias_d, stream),
                (false, None) => ffi::w4a16_load_gptq(self.h, qw, qzeros_p, sc, sd, g_idx_p, bias_p, bias_d, stream),
                (false, Some(n0)) => {
                    ffi::w4a16_load_gptq_part(self.h, n0 as i64, n_part, qw, qzeros_p, sc, sd, g_idx_p, bias_p, bias_d, stream)
                }
            }
        })
    }

    /// Caller-owned output, no allocation, no synchronisation: CUDA-graph capturable (forward_pooled, config.rs:167-174).
    pub fn forward_into(&self, x2: &Tensor, y: &Tensor, m: usize) -> Result<()> {
        check(unsafe { ffi::w4a16_forward_into(self.h, dev_ptr(x2)?, dev_ptr(y)? as *mut c_void, m as i64, stream_ptr(&self.device)?) })
    }

    /// RowParallelLinear::forward (parallel_layers.rs:275-313): fused GEMV + all-reduce over NVLink for m <= 4 when a
    /// peer group is attached, ncclAllReduce otherwise (nccl.rs:1338-1364).
    pub fn forward_row(&self, x2: &Tensor, y: &Tensor, m: usize, peer: *mut ffi::W4a16Peer, comm: *mut c_void) -> Result<()> {
        let (x, yp, st) = (dev_ptr(x2)?, dev_ptr(y)? as *mut c_void, stream_ptr(&self.device)?);
        check(unsafe {
            if peer.is_null() {
                ffi::w4a16_forward_row_allreduce(self.h, x, yp, m as i64, comm, st)
            } else {
                ffi::w4a16_forward_row_fused(self.h, x, yp, m as i64, peer, comm, st)
            }
        })
    }

    /// The on-disk cache of the resident layout: skips the repack of awq_marlin.rs:580-667 on restart.
    pub fn save_tiled(&self, path: &std::path::Path, fingerprint: u64) -> Result<()> {
        let c = std::ffi::CString::new(path.to_string_lossy().as_bytes()).map_err(candle_core::Error::wrap)?;
        check(unsafe { ffi::w4a16_save_tiled(self.h, c.as_ptr(), fingerprint) })
    }
    pub fn load_tiled(&mut self, path: &std::path::Path, fingerprint: u64) -> Result<()> {
        let c = std::ffi::CString::new(path.to_string_lossy().as_bytes()).map_err(candle_core::Error::wrap)?;
        check(unsafe { ffi::w4a16_load_tiled(self.h, c.as_ptr(), fingerprint, stream_ptr(&self.device)?) })
    }
}

impl QuantizedLinear for B200W4A16Linear {
    fn forward(&self, x: &Tensor) -> Result<Tensor> {
        // same contract as awq_marlin.rs:545-571 / marlin_cuda.rs:858-869: BF16, rank >= 2, flatten, restore
        if x.dtype() != DType::BF16 {
            candle_core::bail!("B200W4A16Linear: input dtype must be BF16, got {:?}", x.dtype());
        }
        let dims = x.dims().to_vec();
        if dims.len() < 2 {
            candle_core::bail!("Input must have at least 2 dimensions");
        }
        let k = dims[dims.len() - 1];
        if k != self.in_features() {
            candle_core::bail!("Input K dimension ({k}) doesn't match size_k ({})", self.in_features());
        }
        let m: usize = dims[..dims.len() - 1].iter().product();
        let x2 = x.contiguous()?;
        // every element of y is written by the kernel (marlin_cuda.rs:431-440): no zero-init needed
        let y = unsafe { Tensor::empty((m, self.out_features()), DType::BF16, &self.device)? };
        self.forward_into(&x2, &y, m)?;
        let mut out_dims = dims[..dims.len() - 1].to_vec();
        out_dims.push(self.out_features());
        y.reshape(out_dims)
    }

    fn load_weights(&mut self, w: &HashMap<String, Tensor>) -> Result<()> {
        self.load_impl(w, None)
    }

    fn weight_dtype(&self) -> DType {
        DType::U32
    }
    fn in_features(&self) -> usize {
        unsafe { ffi::w4a16_in_features(self.h) as usize }
    }
    fn out_features(&self) -> usize {
        unsafe { ffi::w4a16_out_features(self.h) as usize }
    }
    fn has_bias(&self) -> bool {
        unsafe { ffi::w4a16_has_bias(self.h) != 0 }
    }
}

impl Drop for B200W4A16Linear {
    fn drop(&mut self) {
        unsafe { ffi::w4a16_destroy(self.h) };
    }
}

/// A fixed sequence of decode linears (M <= 4) as one persistent launch: built once per captured decode
/// graph from the layer loop's (linear, input buffer, output buffer) triples, then `run` per step.  The tensors must
/// stay alive and in place for the life of the chain (they are the graph's arguments).
pub struct DecodeChain {
    c: *mut ffi::W4a16Chain,
    device: Device,
    _keep: Vec<(Tensor, Tensor)>,
}
unsafe impl Send for DecodeChain {}

impl DecodeChain {
    pub fn new(steps: &[(&B200W4A16Linear, Tensor, Tensor)], m: usize) -> Result<Self> {
        let device = steps.first().map(|s| s.0.device.clone()).ok_or_else(|| candle_core::Error::Msg("empty chain".into()))?;
        let mut hs = Vec::with_capacity(steps.len());
        let (mut xs, mut ys) = (Vec::with_capacity(steps.len()), Vec::with_capacity(steps.len()));
        for (lin, x, y) in steps {
            hs.push(lin.raw());
            xs.push(dev_ptr(x)?);
            ys.push(dev_ptr(y)? as *mut c_void);
        }
        let mut c = std::ptr::null_mut();
        check(unsafe { ffi::w4a16_chain_create(&mut c, steps.len() as c_int, hs.as_ptr(), xs.as_ptr(), ys.as_ptr(), m as i64) })?;
        Ok(Self { c, device, _keep: steps.iter().map(|s| (s.1.clone(), s.2.clone())).collect() })
    }
    /// Same, with the residual adds and RMSNorms between the linears done inside the launch (`prologues[i]` for linear i;
    /// `norm_weight` null = none): input_layernorm / post_attention_layernorm + half_add_pooled of
    /// qwen3_quantized.rs:1043-1075.  The tensors behind the raw pointers must outlive the chain.
    pub fn with_prologues(steps: &[(&B200W4A16Linear, Tensor, Tensor)], m: usize, prologues: &[ffi::W4a16ChainPrologue]) -> Result<Self> {
        let device = steps.first().map(|s| s.0.device.clone()).ok_or_else(|| candle_core::Error::Msg("empty chain".into()))?;
        if prologues.len() != steps.len() {
            candle_core::bail!("DecodeChain: one prologue entry per step");
        }
        let hs: Vec<_> = steps.iter().map(|s| s.0.raw()).collect();
        let mut xs = Vec::with_capacity(steps.len());
        let mut ys = Vec::with_capacity(steps.len());
        for (_, x, y) in steps {
            xs.push(dev_ptr(x)?);
            ys.push(dev_ptr(y)? as *mut c_void);
        }
        let mut c = std::ptr::null_mut();
        check(unsafe {
            ffi::w4a16_chain_create_ex(&mut c, steps.len() as c_int, hs.as_ptr(), xs.as_ptr(), ys.as_ptr(), m as i64, prologues.as_ptr())
        })?;
        Ok(Self { c, device, _keep: steps.iter().map(|s| (s.1.clone(), s.2.clone())).collect() })
    }
    /// Tensor parallel: the ROW shards of the chain all-reduce inside the launch over the group's NVLink buffers
    /// (replaces the ncclAllReduce of RowParallelLinear::forward, parallel_layers.rs:295-311).  SPMD: every rank builds
    /// the same chain and runs it the same number of times.
    pub fn set_peer(&self, peer: &PeerGroup) -> Result<()> {
        check(unsafe { ffi::w4a16_chain_set_peer(self.c, peer.raw()) })
    }
    pub fn run(&self) -> Result<()> {
        check(unsafe { ffi::w4a16_chain_run(self.c, stream_ptr(&self.device)?) })
    }
    pub fn kernel(&self) -> String {
        unsafe { CStr::from_ptr(ffi::w4a16_chain_kernel(self.c)) }.to_string_lossy().into_owned()
    }
}
impl Drop for DecodeChain {
    fn drop(&mut self) {
        unsafe { ffi::w4a16_chain_destroy(self.c) };
    }
}

/// Symmetric NVLink buffers of one tensor-parallel group (replaces the dormant kernels/custom_allreduce.cu): created next
/// to `NcclCommunicator::new` (distributed/launcher.rs:204-262); the 64-byte handles travel over the launcher's
/// out-of-band channel like the NCCL unique id.
pub struct PeerGroup {
    p: *mut ffi::W4a16Peer,
}
unsafe impl Send for PeerGroup {}

impl PeerGroup {
    pub fn create(rank: usize, world: usize, max_out_features: usize) -> Result<(Self, [u8; 64])> {
        let mut p = std::ptr::null_mut();
        check(unsafe { ffi::w4a16_peer_create(&mut p, rank as c_int, world as c_int, (max_out_features * 4) as i64) })?;
        let mut handle = [0u8; 64];
        check(unsafe { ffi::w4a16_peer_export(p, handle.as_mut_ptr() as *mut c_void) })?;
        Ok((Self { p }, handle))
    }
    /// `all`: every rank's handle in rank order (world x 64 bytes).
    pub fn attach(&self, all: &[u8]) -> Result<()> {
        check(unsafe { ffi::w4a16_peer_attach(self.p, all.as_ptr() as *const c_void) })
    }
    pub fn raw(&self) -> *mut ffi::W4a16Peer {
        self.p
    }
}
impl Drop for PeerGroup {
    fn drop(&mut self) {
        unsafe { ffi::w4a16_peer_destroy(self.p) };
    }
}

/// check_marlin_supports_shape (marlin.rs:308-341) as the library sees it.
pub fn check_shape(out_features: usize, in_features: usize, group_size: i32) -> Result<()> {
    check(unsafe { ffi::w4a16_check_shape(out_features as i64, in_features as i64, group_size) })
}
