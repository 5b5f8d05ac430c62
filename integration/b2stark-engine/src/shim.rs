//! Glue between openvm-stark-backend's prover traits and the wrappers in lib.rs. The traits live in crates that are not
//! vendored with the reference (openvm-stark-backend @ 002f52e, Plonky3 @ 539bbc84), so the impl blocks are given as a
//! sketch in the module docs below; what IS concrete here is the key-blob writer, whose output format is fixed by
//! include/b2stark.h and exercised by the test-suite through its Python twin (openvm_b200/airs.py::key_blob).
//!
//! ```ignore
//! pub struct B2Backend;   // ProverBackend { Val = BabyBear; Challenge = BinomialExtensionField<BabyBear, 4>;
//!                         //   Commitment = [BabyBear; 8]; Matrix = crate::Matrix; PcsData = crate::PcsData; .. }
//! impl TraceCommitter<B2Backend> for B2Device {
//!     fn commit(&self, traces: &[Matrix]) -> ([BabyBear; 8], PcsData) {
//!         let (root, data) = self.engine.commit(traces).expect("b2s_commit");
//!         (root.map(BabyBear::from_canonical_u32), data)            // vm.rs:735, system/cuda/program.rs:80
//!     }
//! }
//! impl DeviceDataTransporter<SC, B2Backend> for B2Device {
//!     fn transport_matrix_to_device(&self, m: &Arc<RowMajorMatrix<BabyBear>>) -> Matrix {
//!         // BabyBear is a repr(transparent) Montgomery u32: the slice goes down as it is              vm.rs:731-734
//!         let words = unsafe { std::slice::from_raw_parts(m.values.as_ptr() as *const u32, m.values.len()) };
//!         self.engine.upload_row_major(words, m.height() as u64, m.width() as u64, true).expect("upload")
//!     }
//!     fn transport_pk_to_device(&self, pk: &MultiStarkProvingKey<SC>) -> DeviceMultiStarkProvingKey<B2Backend> {
//!         let mut w = KeyBlobWriter::new(pk.get_vk().pre_hash_words(), log_up_pow_bits, pk.per_air.len());
//!         for air in &pk.per_air { w.air(&AirDesc::from(air)); }                                       // vm.rs:382
//!         for c in &pk.trace_height_constraints { w.height_constraint(&c.coefficients, c.threshold); }
//!         self.engine.load_key(&w.finish()).expect("b2s_pk_load")
//!     }
//! }
//! impl StarkEngine<SC> for B2Engine {
//!     fn prove(&self, pk: &DeviceMultiStarkProvingKey<B2Backend>, ctx: ProvingContext<B2Backend>) -> Proof<SC> {
//!         let airs: Vec<AirContext> = ctx.per_air.into_iter().map(|(air_id, c)| AirContext { .. }).collect();
//!         let bytes = self.engine.prove(&pk.inner, &airs).expect("b2s_prove");                         // vm.rs:694,1005
//!         Proof::<SC>::decode(&mut &bytes[..]).expect("codec")                  // crates/sdk/src/codec.rs:332-642
//!     }
//! }
//! ```
//! Selection (crates/sdk/src/lib.rs:71-83):
//! `cfg_if! { if #[cfg(feature = "b2stark")] { pub type DefaultStarkEngine = b2stark_engine::B2Engine; } .. }`

/// Node of a SymbolicExpressionDag as the blob carries it: kind 0 Constant(a) 1 Preprocessed 2 Main 3 Permutation 4 Public
/// 5 Challenge 6 Exposed (a = index, b = offset, c = part) 7 IsFirstRow 8 IsLastRow 9 IsTransition 10 Add 11 Sub 12 Neg
/// 13 Mul (a, b = child indices) -- content consumed at extensions/native/recursion/src/stark/mod.rs:1023-1085.
#[derive(Clone, Copy, Debug)]
pub struct Node {
    pub kind: u32,
    pub a: u32,
    pub b: u32,
    pub c: u32,
}
#[derive(Clone, Debug)]
pub struct Interaction {
    pub bus: u32,
    pub is_receive: bool,
    pub count_node: u32,
    pub field_nodes: Vec<u32>,
}
#[derive(Clone, Debug, Default)]
pub struct Preprocessed {
    pub commit: [u32; 8],
    pub log_height: u32,
    pub row_major_canonical: Vec<u32>,
}
#[derive(Clone, Debug, Default)]
pub struct AirDesc {
    pub width_preprocessed: u32,
    pub cached_widths: Vec<u32>,
    pub width_common_main: u32,
    pub width_after_challenge: u32, // extension-field columns: chunks + 1, or 0
    pub log_quotient_degree: u32,
    pub num_public_values: u32,
    pub preprocessed: Option<Preprocessed>,
    pub nodes: Vec<Node>,
    pub constraint_idx: Vec<u32>,
    pub interactions: Vec<Interaction>,
    pub chunks: Vec<Vec<u32>>, // FRI-LogUp partition: interaction indices per reciprocal-sum column
}

/// Writes the key blob of include/b2stark.h (version 3: trace-height constraints after the AIRs).
pub struct KeyBlobWriter {
    w: Vec<u32>,
    num_airs: usize,
    airs_written: usize,
    constraints: Vec<u32>,
    n_constraints: u32,
}
impl KeyBlobWriter {
    pub fn new(pre_hash: [u32; 8], log_up_pow_bits: u32, num_airs: usize) -> Self {
        let mut w = vec![0x4B50_3242, 3];
        w.extend_from_slice(&pre_hash);
        w.push(log_up_pow_bits);
        w.push(num_airs as u32);
        Self { w, num_airs, airs_written: 0, constraints: Vec::new(), n_constraints: 0 }
    }
    pub fn air(&mut self, a: &AirDesc) {
        assert!(self.airs_written < self.num_airs);
        self.airs_written += 1;
        let w = &mut self.w;
        w.push(a.width_preprocessed);
        w.push(a.cached_widths.len() as u32);
        w.extend_from_slice(&a.cached_widths);
        w.extend_from_slice(&[a.width_common_main, a.width_after_challenge, a.log_quotient_degree, a.num_public_values,
                              if a.width_after_challenge != 0 { 1 } else { 0 }]);
        if a.width_preprocessed != 0 {
            let p = a.preprocessed.as_ref().expect("preprocessed trace");
            w.extend_from_slice(&p.commit);
            w.push(p.log_height);
            assert_eq!(p.row_major_canonical.len(), (a.width_preprocessed as usize) << p.log_height);
            w.extend_from_slice(&p.row_major_canonical);
        }
        w.push(a.nodes.len() as u32);
        for n in &a.nodes {
            w.extend_from_slice(&[n.kind, n.a, n.b, n.c]);
        }
        w.push(a.constraint_idx.len() as u32);
        w.extend_from_slice(&a.constraint_idx);
        w.push(a.interactions.len() as u32);
        for i in &a.interactions {
            w.extend_from_slice(&[i.bus, i.is_receive as u32, i.count_node, i.field_nodes.len() as u32]);
            w.extend_from_slice(&i.field_nodes);
        }
        w.push(a.chunks.len() as u32);
        for c in &a.chunks {
            w.push(c.len() as u32);
            w.extend_from_slice(c);
        }
    }
    /// vk.trace_height_constraints (extensions/native/recursion/src/types.rs:70): one coefficient per AIR, threshold <= p.
    pub fn height_constraint(&mut self, coefficients: &[u32], threshold: u32) {
        assert_eq!(coefficients.len(), self.num_airs);
        self.constraints.extend_from_slice(coefficients);
        self.constraints.push(threshold);
        self.n_constraints += 1;
    }
    pub fn finish(mut self) -> Vec<u32> {
        assert_eq!(self.airs_written, self.num_airs);
        self.w.push(self.n_constraints);
        self.w.extend_from_slice(&self.constraints);
        self.w
    }
}
