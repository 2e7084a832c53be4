//! UNVERIFIED (never compiled) — the `bench` sub-command, written like the sub-commands of `build_app()`
//! (cli/mod.rs:162-232) and read back like `get_command()` reads `deploy` (cli/mod.rs:20-40).
use clap::{App, Arg, ArgMatches, SubCommand};

/// `h2ok bench` as specified by the user (cf. `UserDeploymentSpecification`, cli/mod.rs:88-109).
pub struct BenchSpecification {
    pub shape: String,
    pub rows: u64,
    pub cols: u32,
    pub bins: u32,
    pub depth: u32,
    pub gpus: u32,
    pub trees: u32,
    pub seed: u64,
}

/// Added to the `App` next to "deploy", "undeploy" and "ingress".
pub fn bench_subcommand<'a>() -> App<'a, 'a> {
    SubCommand::with_name("bench")
        .about("Benchmarks the GBM histogram path (histogram build, split search, row routing per tree level) on the GPUs of this node. \
            Prints a table on a terminal, one JSON line otherwise.")
        .arg(Arg::with_name("shape")
            .long("shape")
            .number_of_values(1)
            .default_value("c3")
            .possible_values(&["c2", "c3", "c4"])
            .help("Synthetic data shape: c2 (28 HIGGS-shaped columns), c3 (wide, 64 columns) or c4 (sparse)."))
        .arg(Arg::with_name("rows")
            .long("rows")
            .short("r")
            .number_of_values(1)
            .default_value("0")
            .help("Number of rows over all GPUs. 0: the shape's own (1M / 100M / 10M).")
            .validator(validate_int_not_negative))
        .arg(Arg::with_name("cols")
            .long("cols")
            .number_of_values(1)
            .default_value("0")
            .help("Number of columns. 0: the shape's own (28 / 64 / 2000).")
            .validator(validate_int_not_negative))
        .arg(Arg::with_name("bins")
            .long("bins")
            .short("b")
            .number_of_values(1)
            .default_value("0")
            .help("Histogram bins per column (nbins). 0: the shape's own (256 / 256 / 64).")
            .validator(validate_int_not_negative))
        .arg(Arg::with_name("depth")
            .long("depth")
            .short("d")
            .number_of_values(1)
            .default_value("0")
            .help("Tree depth (levels per tree). 0: the shape's own (6 / 8 / 6).")
            .validator(validate_int_not_negative))
        .arg(Arg::with_name("gpus")
            .long("gpus")
            .short("g")
            .number_of_values(1)
            .default_value("1")
            .help("Number of GPUs of this node to shard the rows over.")
            .validator(super::validate_int_greater_than_zero))
        .arg(Arg::with_name("trees")
            .long("trees")
            .short("t")
            .number_of_values(1)
            .default_value("3")
            .help("Trees to time (one more is built first as a warm-up).")
            .validator(super::validate_int_greater_than_zero))
        .arg(Arg::with_name("seed")
            .long("seed")
            .number_of_values(1)
            .default_value("20261106")
            .help("Seed of the counter-based generator.")
            .validator(validate_int_not_negative))
}

/// Same shape as `validate_int_greater_than_zero` (cli/mod.rs:245-252).
fn validate_int_not_negative(input: String) -> Result<(), String> {
    let number: i64 = input.parse::<i64>().unwrap();
    return if number < 0 {
        Result::Err("Error: The number provided must not be negative.".to_string())
    } else {
        Result::Ok(())
    };
}

/// All arguments have defaults, it is therefore safe to unwrap (as `get_command` does for `deploy`).
pub fn bench_specification(args: &ArgMatches) -> BenchSpecification {
    let shape: String = super::extract_string(args, "shape").unwrap();
    let defaults: (u64, u32, u32, u32) = match shape.as_str() {
        "c2" => (1_000_000, 28, 256, 6),
        "c4" => (10_000_000, 2000, 64, 6),
        _ => (100_000_000, 64, 256, 8),
    };
    let or = |given: u64, default: u64| if given == 0 { default } else { given };
    BenchSpecification {
        rows: or(super::extract_num(args, "rows").unwrap(), defaults.0),
        cols: or(super::extract_num::<u64>(args, "cols").unwrap(), defaults.1 as u64) as u32,
        bins: or(super::extract_num::<u64>(args, "bins").unwrap(), defaults.2 as u64) as u32,
        depth: or(super::extract_num::<u64>(args, "depth").unwrap(), defaults.3 as u64) as u32,
        gpus: super::extract_num(args, "gpus").unwrap(),
        trees: super::extract_num(args, "trees").unwrap(),
        seed: super::extract_num(args, "seed").unwrap(),
        shape,
    }
}
