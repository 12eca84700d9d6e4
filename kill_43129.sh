ssh test@127.0.0.1 "kill -9 5374"
ssh test@127.0.0.1 "kill -9 5375"
ssh test@127.0.0.1 "kill -9 5376"
kill -9 5044
