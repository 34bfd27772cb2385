"""gmlrt - B200-native renderer for the GML `render` operator (drop-in for the hot
path of timdestan/go-raytracer).  See DESIGN.md."""
