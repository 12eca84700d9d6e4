ssh test@127.0.0.1 "kill -9 2198"
ssh test@127.0.0.1 "kill -9 2199"
ssh test@127.0.0.1 "kill -9 2200"
ssh test@127.0.0.1 "kill -9 2201"
kill -9 2109
