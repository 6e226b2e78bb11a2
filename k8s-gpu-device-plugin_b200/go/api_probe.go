// NOT COMPILED in this repository (no Go toolchain in the build image).  The one route added
// to router/api.go:27-37; the four existing routes and the {code,data,msg} envelope
// (modules/util/http.go:3-15) are untouched.  Shape mirrors Restart (router/api.go:50-54).
// host/router.hpp is the tested twin.
package router

import (
	"net/http"

	"github.com/uppercaveman/k8s-gpu-device-plugin/modules/util"

	"github.com/labstack/echo/v4"
)

// in RegistApiRouter:  root.GET("/probe", a.Probe)

// Probe : 按需重新探测 — re-grade now (all devices, or ?uuid=GPU-…), feed ListAndWatch and
// the gauges, answer with the results.
func (a *API) Probe(c echo.Context) error {
	results, err := a.pluginManager.Reprobe(c.QueryParam("uuid"), false) // on demand: busy devices are probed too
	if err != nil {
		return c.JSON(http.StatusOK, util.Failed(404, err.Error()))
	}
	return c.JSON(http.StatusOK, util.Success(results))
}
