//! `B200Kernel`: the `KernelProvider` BitNet-rs would register for an sm_100 device. It is the reference's own
//! `FfiKernel` (crates/bitnet-kernels/src/ffi/bridge.rs:159-263) pointed at libbitnet_b200.so: same refcounted
//! init/cleanup, same slice-length checks and messages, same `rc -> KernelError::ExecutionFailed` mapping.
//!
//! Registration (crates/bitnet-kernels/src/lib.rs:67-120, `KernelManager::new`):
//! ```ignore
//! if let Ok(k) = bitnet_b200_sys::b200::B200Kernel::new() { providers.insert(0, Box::new(k)); }
//! ```
//! and `select_gpu_kernel` (lib.rs:243-250) returns it for CUDA device ordinals.
use crate::*;
use bitnet_common::{BitNetError, KernelError, QuantizationType, Result};
use bitnet_kernels::KernelProvider;
use libc::c_int;

pub struct B200Kernel {
    _initialized: bool,
}

fn exec_failed(reason: String) -> BitNetError {
    BitNetError::Kernel(KernelError::ExecutionFailed { reason })
}

fn rc(code: c_int) -> Result<()> {
    // bridge.rs:223-227
    if code == 0 {
        return Ok(());
    }
    Err(exec_failed(format!("B200 kernel error: {}", last_error())))
}

impl B200Kernel {
    /// bridge.rs:165-174: `bitnet_cpp_init` is reference-counted on the C side; no CPU fallback is installed
    /// when it fails (north_star), the error is the caller's to handle.
    pub fn new() -> Result<Self> {
        // SAFETY: plain C call without arguments.
        if unsafe { bitnet_cpp_init() } != 0 {
            return Err(exec_failed(format!("B200 init failed: {}", last_error())));
        }
        Ok(Self { _initialized: true })
    }

    /// The live-path seam (crates/bitnet-quantization/src/i2s_qk256.rs:346-372), host slices.
    pub fn gemv_qk256(
        qs: &[u8], x: &[f32], y: &mut [f32], rows: usize, cols: usize, row_stride_bytes: usize,
    ) -> anyhow::Result<()> {
        // SAFETY: pointers and lengths come from live slices; the callee validates them against rows/cols.
        let code = unsafe {
            bitnet_b200_gemv_qk256(
                qs.as_ptr(), qs.len(), x.as_ptr(), x.len(), y.as_mut_ptr(), y.len(), rows, cols, row_stride_bytes,
            )
        };
        if code != 0 {
            anyhow::bail!("{}", last_error()); // carries "y_out length" / "x length" / "too short"
        }
        Ok(())
    }
}

impl Drop for B200Kernel {
    fn drop(&mut self) {
        // SAFETY: balanced with the successful init in `new`.
        unsafe { bitnet_cpp_cleanup() }
    }
}

impl KernelProvider for B200Kernel {
    fn name(&self) -> &'static str {
        "b200"
    }

    fn is_available(&self) -> bool {
        // SAFETY: plain C call without arguments.
        unsafe { bitnet_cpp_is_available() != 0 }
    }

    fn matmul_i2s(&self, a: &[i8], b: &[u8], c: &mut [f32], m: usize, n: usize, k: usize) -> Result<()> {
        // bridge.rs:205-220 / cpu/fallback.rs:48-63
        if a.len() != m * k {
            return Err(exec_failed(format!("Matrix A dimension mismatch: expected {}, got {}", m * k, a.len())));
        }
        if b.len() != k * n {
            return Err(exec_failed(format!("Matrix B dimension mismatch: expected {}, got {}", k * n, b.len())));
        }
        if c.len() != m * n {
            return Err(exec_failed(format!("Matrix C dimension mismatch: expected {}, got {}", m * n, c.len())));
        }
        // SAFETY: lengths checked above.
        rc(unsafe { bitnet_cpp_matmul_i2s(a.as_ptr(), b.as_ptr(), c.as_mut_ptr(), m as c_int, n as c_int, k as c_int) })
    }

    fn quantize(&self, input: &[f32], output: &mut [u8], scales: &mut [f32], qtype: QuantizationType) -> Result<()> {
        // bridge.rs:229-262: qtype travels as its discriminant (I2S = 0, TL1 = 1, TL2 = 2)
        // SAFETY: pointers and lengths come from live slices; the callee checks the output sizes.
        rc(unsafe {
            bitnet_cpp_quantize(
                input.as_ptr(), input.len() as c_int, output.as_mut_ptr(), output.len() as c_int, scales.as_mut_ptr(),
                scales.len() as c_int, qtype as c_int,
            )
        })
    }
}
