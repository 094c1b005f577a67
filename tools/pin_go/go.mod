module pin_go

go 1.23.0

// the reference's pins (/root/reference/go.mod:7,35,36)
require (
	github.com/prometheus/client_golang v1.20.2
	github.com/prometheus/client_model v0.6.1 // indirect
	github.com/prometheus/common v0.55.0 // indirect
)
