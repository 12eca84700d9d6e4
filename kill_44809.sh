ssh test@127.0.0.1 "kill -9 2554"
ssh test@127.0.0.1 "kill -9 2555"
ssh test@127.0.0.1 "kill -9 2556"
ssh test@127.0.0.1 "kill -9 2557"
kill -9 2109
