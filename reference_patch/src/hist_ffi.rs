//! UNVERIFIED (never compiled: no Rust toolchain on the build machine) — binding of libh2ok_hist.so for `h2ok bench`.
//!
//! Only the entry points the sub-command needs; the full C ABI is include/h2ok_hist.h.  The library drives every GPU
//! of the group from the calling thread, which keeps `h2ok` single-threaded and blocking like the rest of the crate
//! (main.rs:18-37, k8s/mod.rs:85-93).  Every call returns a status; the message comes from `h2ok_last_error()`.
use std::ffi::CStr;
use std::os::raw::{c_char, c_float, c_int};

#[repr(C)]
pub struct H2okGroup {
    _private: [u8; 0],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct H2okParams {
    pub nbins: i32,
    pub nbins_top_level: i32,
    pub max_depth: i32,
    pub pad: i32,
    pub min_rows: f64,
    pub min_split_improvement: f64,
}

/// `h2ok_split` (ABI version 2, 208 bytes): DTree.Split as chosen for one node.  `equal == 2`: a categorical split,
/// category c goes right when bit c of `cat_bits` is set; otherwise `x >= splat` goes right.  NaN follows `na_dir`.
#[repr(C)]
#[derive(Clone, Copy)]
pub struct H2okSplit {
    pub col: i32,
    pub bin: i32,
    pub na_dir: i32,
    pub splat: f32,
    pub se_before: f64,
    pub se_left: f64,
    pub se_right: f64,
    pub n_left: f64,
    pub n_right: f64,
    pub pred_left: f64,
    pub pred_right: f64,
    pub equal: i32,
    pub pad: i32,
    pub cat_bits: [u32; 32],
}

/// `h2ok_level_out`: host buffers `h2ok_hist_group_level` fills for the level it ran (any pointer may be null).
#[repr(C)]
pub struct H2okLevelOut {
    pub hist: *mut f64,
    pub min2: *mut f32,
    pub maxin: *mut f32,
    pub splits: *mut H2okSplit,
    pub child_descs: *mut u8, // h2ok_col_desc[n_children][cols], 32 bytes each
    pub child_ids: *mut i32,
}

pub const H2OK_SHAPE_HIGGS28: i32 = 2;
pub const H2OK_SHAPE_WIDE64: i32 = 3;
pub const H2OK_SHAPE_SPARSE: i32 = 4;
pub const H2OK_REDUCE_SCATTER: i32 = 0;

extern "C" {
    fn h2ok_last_error() -> *const c_char;
    fn h2ok_hist_group_create(device_ids: *const i32, n_dev: i32, out: *mut *mut H2okGroup) -> i32;
    fn h2ok_hist_group_destroy(group: *mut H2okGroup) -> i32;
    fn h2ok_hist_group_load_synth(group: *mut H2okGroup, shape: i32, seed: u64, rows_total: i64, cols: i32, na_permille: i32) -> i32;
    fn h2ok_hist_group_loop_create(group: *mut H2okGroup, params: *const H2okParams, is_int: *const i32) -> i32;
    fn h2ok_hist_group_set_options(group: *mut H2okGroup, reduce_mode: i32, checksum: i32, time_phases: i32) -> i32;
    fn h2ok_hist_group_begin_tree_synth(group: *mut H2okGroup, seed: u64, weight_mode: i32) -> i32;
    fn h2ok_hist_group_level(group: *mut H2okGroup, out: *const H2okLevelOut, n_nodes_done: *mut i32, n_children: *mut i32) -> i32;
    fn h2ok_hist_group_checksum(group: *mut H2okGroup, out3: *mut u64) -> i32;
    fn h2ok_hist_group_hist_ms(group: *mut H2okGroup, ms_out: *mut c_float) -> i32;
    fn h2ok_hist_group_sync(group: *mut H2okGroup) -> i32;
}

fn check(status: c_int) -> Result<(), String> {
    if status == 0 {
        return Ok(());
    }
    let msg = unsafe { CStr::from_ptr(h2ok_last_error()) }.to_string_lossy().into_owned();
    Err(format!("h2ok_hist status {}: {}", status, msg))
}

/// What `h2ok bench` reports.
pub struct BenchReport {
    pub ms_per_level: f64,
    pub hist_ms_per_level: f64,
    pub cells_per_second: f64,
    pub levels: u64,
    pub checksum: [u64; 3],
}

/// Owns the group; destroyed on drop so an early `?` cannot leak device memory.
struct Group(*mut H2okGroup);

impl Drop for Group {
    fn drop(&mut self) {
        unsafe { h2ok_hist_group_destroy(self.0) };
    }
}

pub fn run_bench(shape: i32, rows: i64, cols: i32, bins: i32, depth: i32, gpus: u32, trees: u32, seed: u64) -> Result<BenchReport, String> {
    let devices: Vec<i32> = (0..gpus as i32).collect();
    let mut raw: *mut H2okGroup = std::ptr::null_mut();
    check(unsafe { h2ok_hist_group_create(devices.as_ptr(), devices.len() as i32, &mut raw) })?;
    let group = Group(raw);
    check(unsafe { h2ok_hist_group_load_synth(group.0, shape, seed, rows, cols, 0) })?;
    let params = H2okParams { nbins: bins, nbins_top_level: bins, max_depth: depth, pad: 0, min_rows: 10.0, min_split_improvement: 1e-5 };
    check(unsafe { h2ok_hist_group_loop_create(group.0, &params, std::ptr::null()) })?;
    check(unsafe { h2ok_hist_group_set_options(group.0, H2OK_REDUCE_SCATTER, 1, 0) })?;

    let (mut seconds, mut hist_ms, mut levels) = (0.0f64, 0.0f64, 0u64);
    let mut checksum = [0u64; 3];
    for tree in 0..=trees {
        // tree 0 warms up
        check(unsafe { h2ok_hist_group_begin_tree_synth(group.0, seed, 0) })?;
        check(unsafe { h2ok_hist_group_sync(group.0) })?;
        let started = std::time::Instant::now();
        let (mut children, mut level, mut tree_hist_ms) = (1i32, 0i32, 0.0f64);
        while children > 0 && level < depth {
            check(unsafe { h2ok_hist_group_level(group.0, std::ptr::null(), std::ptr::null_mut(), &mut children) })?;
            let mut ms: c_float = 0.0;
            check(unsafe { h2ok_hist_group_hist_ms(group.0, &mut ms) })?;
            tree_hist_ms += ms as f64;
            level += 1;
        }
        check(unsafe { h2ok_hist_group_sync(group.0) })?;
        if tree > 0 {
            seconds += started.elapsed().as_secs_f64();
            hist_ms += tree_hist_ms;
            levels += level as u64;
        }
        if tree == trees {
            check(unsafe { h2ok_hist_group_checksum(group.0, checksum.as_mut_ptr()) })?;
        }
    }
    Ok(BenchReport {
        ms_per_level: seconds * 1e3 / levels as f64,
        hist_ms_per_level: hist_ms / levels as f64,
        cells_per_second: rows as f64 * cols as f64 * levels as f64 / seconds,
        levels,
        checksum,
    })
}
