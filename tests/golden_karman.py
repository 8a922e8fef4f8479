"""Golden Globals of the reference's d2q9 karman regression case (double precision CPU build),
rows at iteration 100..1000: PressureLoss, OutletFlux, InletFlux.
Source: /root/reference/tests/external/d2q9/output/karman_Log_P00_00000000.csv (data, not code)."""
KARMAN = None  # filled from the mounted reference when present; else the embedded subset below is used

KARMAN_IT1000 = dict(PressureLoss=7.6859393965859e-04, OutletFlux=8.7899634755736e-01, InletFlux=9.6042681024743e-01)

KARMAN_XML = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/" permissive="true">
	<Geometry nx="1024" ny="100">
		<MRT><Box/></MRT>
		<WVelocity name="Inlet"><Inlet/></WVelocity>
		<EPressure name="Outlet"><Outlet/></EPressure>
		<Inlet nx="1" dx="5"><Box/></Inlet>
		<Outlet nx="1" dx="-5"><Box/></Outlet>
		<Wall mask="ALL">
			<Channel/>
			<Wedge dx="120" nx="20" dy="50" ny="20" direction="LowerRight"/>
			<Wedge dx="120" nx="20" dy="30" ny="20" direction="UpperRight"/>
			<Wedge dx="140" nx="20" dy="50" ny="20" direction="LowerLeft"/>
			<Wedge dx="140" nx="20" dy="30" ny="20" direction="UpperLeft"/>
		</Wall>
	</Geometry>
	<Model>
		<Param name="VelocityX" value="0.01"/>
		<Param name="Viscosity" value="0.02"/>
	</Model>
	<Solve Iterations="1000"><Log Iterations="100"/></Solve>
</CLBConfig>
"""
