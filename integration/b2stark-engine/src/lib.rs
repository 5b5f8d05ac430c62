//! Owning wrappers over libb2stark.so (include/b2stark.h): engine, device matrices, commitments, keys, proofs.
//! NOT COMPILED in this repository's build environment (no Rust toolchain); the C ABI underneath is what the test-suite
//! exercises through ctypes (openvm_b200/_lib.py) with the same calls in the same order.
//!
//! Reference interface each type stands in for:
//!   Engine      E: StarkFriEngine            crates/sdk/src/prover/vm/mod.rs:22, crates/vm/src/utils/stark_utils.rs:115
//!   Matrix      DeviceMatrix<F>              crates/vm/src/system/cuda/program.rs:80,119
//!   PcsData     CommittedTraceData.data      crates/vm/src/arch/vm.rs:748-763
//!   ProvingKey  DeviceMultiStarkProvingKey   crates/vm/src/arch/vm.rs:382
pub mod ffi;
#[cfg(feature = "openvm")]
pub mod shim;

use std::{ffi::CStr, ptr, sync::Arc};

#[derive(Debug, Clone)]
pub struct Error {
    pub status: i32,
    pub message: String,
}
pub type Result<T> = std::result::Result<T, Error>;

pub use ffi::B2sFriParams as FriParams;

struct EngineHandle(*mut ffi::B2sEngine);
unsafe impl Send for EngineHandle {}
impl Drop for EngineHandle {
    fn drop(&mut self) {
        unsafe { ffi::b2s_engine_free(self.0) }
    }
}

/// One engine = one CUDA stream on one device; used from one thread at a time (`&mut self` in vm.prove, vm.rs:671-697).
#[derive(Clone)]
pub struct Engine {
    h: Arc<EngineHandle>,
    params: FriParams,
}

pub const HASH_POSEIDON2_BABYBEAR: i32 = ffi::B2S_HASH_POSEIDON2_BABYBEAR; // BabyBearPoseidon2Config
pub const HASH_POSEIDON2_BN254: i32 = ffi::B2S_HASH_POSEIDON2_BN254; // BabyBearPoseidon2RootConfig

impl Engine {
    pub fn new(params: FriParams, cuda_device: i32) -> Result<Self> {
        let mut h = ptr::null_mut();
        let st = unsafe { ffi::b2s_engine_new(&params, cuda_device, &mut h) };
        if st != 0 {
            return Err(last_error(ptr::null(), st));
        }
        Ok(Self { h: Arc::new(EngineHandle(h)), params })
    }
    pub fn fri_params(&self) -> FriParams {
        self.params
    }
    fn raw(&self) -> *mut ffi::B2sEngine {
        self.h.0
    }
    fn check(&self, st: i32) -> Result<()> {
        if st == 0 { Ok(()) } else { Err(last_error(self.raw(), st)) }
    }
    /// Keys must be loaded after this call: preprocessed commitments are digests of the selected hash.
    pub fn set_hash(&self, kind: i32) -> Result<()> {
        self.check(unsafe { ffi::b2s_engine_set_hash(self.raw(), kind) })
    }
    /// Share the stream the in-tree CUDA tracegen launches on.
    pub fn set_stream(&self, cuda_stream: *mut std::os::raw::c_void) -> Result<()> {
        self.check(unsafe { ffi::b2s_engine_set_stream(self.raw(), cuda_stream) })
    }
    pub fn metrics(&self) -> Result<ffi::B2sMetrics> {
        let mut m = unsafe { std::mem::zeroed() };
        self.check(unsafe { ffi::b2s_get_metrics(self.raw(), &mut m) })?;
        Ok(m)
    }
    /// The reference's stage gauges (crates/prof/data/metrics.example.json, emitted around crates/vm/src/arch/vm.rs:1328-1399)
    /// as (name, milliseconds) pairs: the first six are disjoint intervals of the proofs since the last reset, the seventh is
    /// their total. A shim forwards each pair to `metrics::gauge!(name).set(ms)` where the reference emits it.
    pub fn stage_gauges(&self) -> Result<[(&'static str, f64); 7]> {
        let m = self.metrics()?;
        Ok([
            ("main_trace_commit_time_ms", m.main_trace_commit_time_ms),
            ("generate_perm_trace_time_ms", m.generate_perm_trace_time_ms),
            ("perm_trace_commit_time_ms", m.perm_trace_commit_time_ms),
            ("quotient_poly_compute_time_ms", m.quotient_poly_compute_time_ms),
            ("quotient_poly_commit_time_ms", m.quotient_poly_commit_time_ms),
            ("pcs_opening_time_ms", m.pcs_opening_time_ms),
            ("stark_prove_excluding_trace_time_ms", m.stark_prove_excluding_trace_time_ms),
        ])
    }
    pub fn reset_metrics(&self) -> Result<()> {
        self.check(unsafe { ffi::b2s_reset_metrics(self.raw()) })
    }

    /// transport_matrix_to_device: `words` is a RowMajorMatrix<BabyBear> viewed as u32 (Plonky3's BabyBear is a
    /// repr(transparent) Montgomery word: pass `monty = true`) or canonical values.
    pub fn upload_row_major(&self, words: &[u32], height: u64, width: u64, monty: bool) -> Result<Matrix> {
        assert_eq!(words.len() as u64, height * width);
        let mut m = ptr::null_mut();
        let repr = if monty { ffi::B2S_REPR_MONTY } else { ffi::B2S_REPR_CANONICAL };
        self.check(unsafe { ffi::b2s_matrix_upload(self.raw(), words.as_ptr(), height, width, ffi::B2S_LAYOUT_ROW_MAJOR, repr, &mut m) })?;
        Ok(Matrix(Arc::new(MatrixHandle { m, _eng: self.h.clone() })))
    }
    /// Zero-copy adoption of a column-major Montgomery device buffer written by the CUDA tracegen (docs/repo/cuda.md).
    /// The buffer must outlive the returned matrix.
    pub unsafe fn adopt_device(&self, device_ptr: *mut std::os::raw::c_void, height: u64, width: u64) -> Result<Matrix> {
        let mut m = ptr::null_mut();
        self.check(ffi::b2s_matrix_adopt_device(self.raw(), device_ptr, height, width, &mut m))?;
        Ok(Matrix(Arc::new(MatrixHandle { m, _eng: self.h.clone() })))
    }
    /// TraceCommitter::commit: root (8 canonical BabyBear words, or 8 limbs of one BN254 scalar) and the prover data.
    pub fn commit(&self, traces: &[Matrix]) -> Result<([u32; 8], PcsData)> {
        let ptrs: Vec<*const ffi::B2sMatrix> = traces.iter().map(|t| t.0.m as *const _).collect();
        let mut root = [0u32; 8];
        let mut d = ptr::null_mut();
        self.check(unsafe { ffi::b2s_commit(self.raw(), ptrs.as_ptr(), ptrs.len(), root.as_mut_ptr(), &mut d) })?;
        Ok((root, PcsData(Arc::new(PcsHandle { d, _eng: self.h.clone() }))))
    }
    /// transport_pk_to_device: the key blob documented in include/b2stark.h (shim.rs flattens a MultiStarkProvingKey).
    pub fn load_key(&self, blob: &[u32]) -> Result<ProvingKey> {
        let mut pk = ptr::null_mut();
        self.check(unsafe { ffi::b2s_pk_load(self.raw(), blob.as_ptr(), blob.len(), &mut pk) })?;
        Ok(ProvingKey { pk, _eng: self.h.clone() })
    }
    /// StarkEngine::prove: proof bytes in the wire format of crates/sdk/src/codec.rs.
    pub fn prove(&self, pk: &ProvingKey, airs: &[AirContext]) -> Result<Vec<u8>> {
        let cached: Vec<Vec<*const ffi::B2sPcsData>> = airs.iter().map(|a| a.cached_mains.iter().map(|c| c.0.d as *const _).collect()).collect();
        let traces: Vec<Vec<*const ffi::B2sMatrix>> = airs.iter().map(|a| a.cached_traces.iter().map(|t| t.0.m as *const _).collect()).collect();
        let ctx: Vec<ffi::B2sAirCtx> = airs
            .iter()
            .enumerate()
            .map(|(i, a)| ffi::B2sAirCtx {
                air_id: a.air_id,
                n_cached: cached[i].len() as u32,
                cached: if cached[i].is_empty() { ptr::null() } else { cached[i].as_ptr() },
                cached_traces: if traces[i].is_empty() { ptr::null() } else { traces[i].as_ptr() },
                common_main: a.common_main.as_ref().map_or(ptr::null(), |m| m.0.m as *const _),
                public_values: if a.public_values.is_empty() { ptr::null() } else { a.public_values.as_ptr() },
                n_public_values: a.public_values.len() as u32,
            })
            .collect();
        let (mut bytes, mut len) = (ptr::null_mut(), 0usize);
        self.check(unsafe { ffi::b2s_prove(self.raw(), pk.pk, ctx.as_ptr(), ctx.len(), &mut bytes, &mut len) })?;
        let out = unsafe { std::slice::from_raw_parts(bytes, len) }.to_vec();
        unsafe { ffi::b2s_free(bytes as *mut _) };
        Ok(out)
    }
    /// StarkEngine::verify: Err carries the reference's VerificationError variant name at the start of the message.
    pub fn verify(&self, pk: &ProvingKey, proof: &[u8]) -> Result<()> {
        self.check(unsafe { ffi::b2s_verify(self.raw(), pk.pk, proof.as_ptr(), proof.len()) })
    }
}

/// proof.write() of the next aggregation layer's verifier program (recursion/src/hints.rs:272-316): vectors of words.
/// Merkle root of one already-extended device matrix with NO allocation (digest layers in a static device array): for callers
/// under CUDA graph capture or in loops that must not call cudaMalloc. At most `B2S_STATIC_MAX_ROWS` rows; calls that overlap
/// on the GPU race on the shared scratch. `d_root8` is device memory for 8 canonical words. No counterpart in the reference.
pub unsafe fn merkle_root_static(d_matrix: *const u32, height: u64, width: u64, d_root8: *mut u32,
                                 cuda_stream: *mut std::os::raw::c_void) -> std::result::Result<(), i32> {
    match ffi::b2s_merkle_root_static(d_matrix, height, width, d_root8, cuda_stream) {
        0 => Ok(()),
        st => Err(st),
    }
}

pub fn proof_hint_stream(proof: &[u8]) -> Result<Vec<Vec<u32>>> {
    let (mut words, mut n) = (ptr::null_mut(), 0usize);
    let st = unsafe { ffi::b2s_proof_hint_stream(ptr::null_mut(), proof.as_ptr(), proof.len(), &mut words, &mut n) };
    if st != 0 {
        return Err(last_error(ptr::null(), st));
    }
    let flat = unsafe { std::slice::from_raw_parts(words, n) }.to_vec();
    unsafe { ffi::b2s_free(words as *mut _) };
    let mut out = Vec::with_capacity(flat[0] as usize);
    let mut i = 1;
    for _ in 0..flat[0] {
        let len = flat[i] as usize;
        out.push(flat[i + 1..i + 1 + len].to_vec());
        i += 1 + len;
    }
    Ok(out)
}

pub struct AirContext {
    pub air_id: u32,
    pub common_main: Option<Matrix>,
    pub public_values: Vec<u32>, // canonical
    pub cached_mains: Vec<PcsData>,
    pub cached_traces: Vec<Matrix>, // needed by AIRs with interactions
}

struct MatrixHandle {
    m: *mut ffi::B2sMatrix,
    _eng: Arc<EngineHandle>,
}
impl Drop for MatrixHandle {
    fn drop(&mut self) {
        unsafe { ffi::b2s_matrix_free(self.m) }
    }
}
/// Ref-counted like the reference's DeviceMatrix (Arc<DeviceBuffer>): cloned freely, freed with the last clone.
#[derive(Clone)]
pub struct Matrix(Arc<MatrixHandle>);
impl Matrix {
    pub fn height(&self) -> u64 {
        unsafe { ffi::b2s_matrix_height(self.0.m) }
    }
    pub fn width(&self) -> u64 {
        unsafe { ffi::b2s_matrix_width(self.0.m) }
    }
}

struct PcsHandle {
    d: *mut ffi::B2sPcsData,
    _eng: Arc<EngineHandle>,
}
impl Drop for PcsHandle {
    fn drop(&mut self) {
        unsafe { ffi::b2s_pcs_free(self.d) }
    }
}
/// Cached across proofs like CommittedTraceData (vm.rs:761-763).
#[derive(Clone)]
pub struct PcsData(Arc<PcsHandle>);

pub struct ProvingKey {
    pk: *mut ffi::B2sPk,
    _eng: Arc<EngineHandle>,
}
impl Drop for ProvingKey {
    fn drop(&mut self) {
        unsafe { ffi::b2s_pk_free(self.pk) }
    }
}

fn last_error(e: *const ffi::B2sEngine, status: i32) -> Error {
    let p = unsafe { ffi::b2s_last_error(e) };
    let message = if p.is_null() { String::new() } else { unsafe { CStr::from_ptr(p) }.to_string_lossy().into_owned() };
    Error { status, message }
}
