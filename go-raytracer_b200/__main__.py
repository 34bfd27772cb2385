"""`python -m go_raytracer_b200 --gml_file scene.gml [--out_file out.png|out.ppm]` - the twin of the reference's
cmd/example (cmd/example/main.go:14-62): parse + evaluate a GML file, render it on the GPU through the C ABI and write
the image.  `.ppm` writes the binary P6 the ICFP task (and the GML `render` operator's file argument) names; anything
else writes PNG like the reference.  No CPU fallback: without a CUDA device this exits with the library's error."""
import argparse
import os
import sys

from . import cuda, imageio, raytracer


def main(argv=None):
    ap = argparse.ArgumentParser(prog="go_raytracer_b200")
    ap.add_argument("--gml_file", required=True, help="gml filename to run")
    ap.add_argument("--out_file", default="", help="png / ppm filename to write (default: output/<name>.png)")
    ap.add_argument("--precision", default="f32", choices=["f32", "f64"], help="f64 = bit-parity mode")
    ap.add_argument("--devices", default="0", help="comma-separated CUDA ordinals")
    args = ap.parse_args(argv)
    out = args.out_file
    if not out:
        base = os.path.basename(args.gml_file)
        if not base.endswith(".gml"):
            sys.exit("Could not derive --out_file, please specify it.")
        out = os.path.join("output", base[:-4] + ".png")
        print(f"Using derived output path: {out}", file=sys.stderr)
    prec = cuda.PRECISION_F32 if args.precision == "f32" else cuda.PRECISION_F64_PARITY
    img = raytracer.ParseAndRenderGMLFile(args.gml_file, devices=tuple(int(d) for d in args.devices.split(",")), precision=prec)
    os.makedirs(os.path.dirname(os.path.abspath(out)), exist_ok=True)
    (imageio.write_ppm if out.lower().endswith(".ppm") else imageio.write_png)(out, img)
    print(f"wrote {out}")


if __name__ == "__main__":
    main()
